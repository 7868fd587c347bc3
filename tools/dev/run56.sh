ONILDA_TIMING=1 ONILDA_PDL=0 timeout 300 python tools/estep_sweep.py --iters 21 "" 2>&1 | grep "^\[onilda\]" | awk 'NR>=4' | head -20
