ONILDA_TIMING=1 timeout 300 python tools/estep_sweep.py --iters 6 "" 2>&1 | grep "^\[onilda\]" | tail -3
timeout 1200 python tools/estep_sweep.py --iters 21 "ONILDA_LONG_LEN=768" "ONILDA_LONG_LEN=512" "ONILDA_LONG_LEN=384" "ONILDA_LONG_LEN=512 ONILDA_LONG_PER_SM=2" "ONILDA_LONG_LEN=256 ONILDA_LONG_PER_SM=2" "ONILDA_LONG_THREADS=128 ONILDA_LONG_PER_SM=2" 2>&1 | grep "estep_ms\|rror\|FAIL"
