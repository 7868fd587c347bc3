ONILDA_SWEEP_COUNTS=1 timeout 300 python tools/estep_sweep.py --iters 3 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" 2>&1 | grep "estep_ms\|FAILED\|rror"
ONILDA_SWEEP_COUNTS=1 timeout 300 python tools/estep_sweep.py --iters 12 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" 2>&1 | grep "estep_ms\|FAILED\|rror"
