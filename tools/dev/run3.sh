python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4
ONILDA_TIMING=1 timeout 600 python tools/estep_sweep.py --iters 3 --num-docs 592 --hub-frac 1.0 \
  "ONILDA_LONG_THREADS=128" "ONILDA_LONG_THREADS=128 ONILDA_LONG_GROUP=16" > gpurun_out/r2_hub592b.log 2>&1
grep -v "^\[onilda\]" gpurun_out/r2_hub592b.log
ONILDA_TIMING=1 timeout 900 python tools/estep_sweep.py --iters 6 \
  "ONILDA_BULK_THREADS=128" "ONILDA_BULK_THREADS=128 ONILDA_LONG_GROUP=16" "ONILDA_BULK_THREADS=192" \
  "ONILDA_BULK_THREADS=128 ONILDA_LONG_THREADS=128 ONILDA_LONG_PER_SM=2" \
  > gpurun_out/r2_sweep3.log 2>&1
awk '/^\[onilda\]/{last=$0} /estep_ms|corpus/{print substr($0,1,150); print "    " last}' gpurun_out/r2_sweep3.log
