timeout 1200 python tools/estep_sweep.py --iters 12 "" "ONILDA_LONG_ACTIVE=7" "ONILDA_LONG_ACTIVE=6" "ONILDA_LONG_ACTIVE=5" "ONILDA_LONG_ACTIVE=4" 2>&1 | grep "estep_ms\|rror\|FAIL"
