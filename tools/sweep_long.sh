run() { echo "== $*"; env "$@" ONILDA_TIMING=1 python bench.py --steps 4 --warmup 3 --no-e2e --no-scoring --no-grouping --no-cpu-baseline 2>gpurun_out/sw_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['em']['estep_ms'])"; grep -m2 "long\|bulk" gpurun_out/sw_err.log | head -2; }
run A=1
run ONILDA_LONG_GROUP=8
run ONILDA_LONG_GROUP=8 ONILDA_LONG_SMEM=57184
run ONILDA_LONG_GROUP=16
run ONILDA_LONG_GROUP=16 ONILDA_LONG_SMEM=57184
run ONILDA_LONG_GROUP=8 ONILDA_LONG_LEN=4096
