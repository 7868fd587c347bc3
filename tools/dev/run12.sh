timeout 900 python tools/estep_sweep.py --iters 21 "ONILDA_BULK_THREADS=128" "ONILDA_BULK_THREADS=256" 2>&1 | grep estep_ms
