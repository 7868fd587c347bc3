python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
ONILDA_TIMING=1 timeout 900 python tools/estep_sweep.py --iters 21 "" "" "ONILDA_BULK_THREADS=128" > gpurun_out/r2_sweep25.log 2>&1
grep estep_ms gpurun_out/r2_sweep25.log
grep "E-step 1[3-9][0-9]\|E-step 2" gpurun_out/r2_sweep25.log | head -12
