timeout 900 python tools/estep_sweep.py --iters 21 "ONILDA_BULK_WAVES=1" "ONILDA_BULK_WAVES=2" "ONILDA_BULK_WAVES=1" "ONILDA_BULK_WAVES=1 ONILDA_BULK_THREADS=128" 2>&1 | grep estep_ms
