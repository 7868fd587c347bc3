python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/estep_sweep.py --iters 3 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" "ONILDA_LONG_THREADS=256" 2>&1 | grep "estep_ms\|FAILED\|rror"
ONILDA_TIMING=1 timeout 900 python tools/estep_sweep.py --iters 6 \
  "ONILDA_BULK_THREADS=128" "ONILDA_BULK_THREADS=128 ONILDA_LONG_THREADS=128" \
  > gpurun_out/r2_sweep5.log 2>&1
awk '/^\[onilda\]/{last=$0} /estep_ms|corpus/{print substr($0,1,150); print "    " last}' gpurun_out/r2_sweep5.log
timeout 900 ncu --set full --import-source on --clock-control none -k regex:estep_long -c 1 --launch-skip 1 \
  -o gpurun_out/r2_long_alone3 python tools/estep_sweep.py --iters 2 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" > gpurun_out/r2_ncu_long3.log 2>&1
tail -2 gpurun_out/r2_ncu_long3.log
