nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/l2_gather tools/l2_gather_bench.cu && /tmp/l2_gather | tee gpurun_out/r2_l2_gather.txt
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
ONILDA_SCORE_L2_PERSIST=0 python tools/score_time.py 2>&1 | tail -1
ONILDA_SCORE_L2_PERSIST=1 python tools/score_time.py 2>&1 | tail -1
