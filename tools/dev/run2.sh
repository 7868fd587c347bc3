set -x
ONILDA_TIMING=1 timeout 600 python tools/estep_sweep.py --iters 4 --num-docs 592 --hub-frac 1.0 \
  "ONILDA_LONG_THREADS=128" "ONILDA_LONG_THREADS=128 ONILDA_LONG_GROUP=16" "ONILDA_LONG_THREADS=128 ONILDA_LONG_GROUP=8" \
  "ONILDA_LONG_THREADS=128 ONILDA_LONG_GROUP=4" > gpurun_out/r2_hub592.log 2>&1
grep -v "^\[onilda\]" gpurun_out/r2_hub592.log
ONILDA_TIMING=1 timeout 600 python tools/estep_sweep.py --iters 4 --num-docs 1184 --hub-frac 1.0 \
  "ONILDA_LONG_THREADS=256" "ONILDA_LONG_THREADS=256 ONILDA_LONG_GROUP=16" "ONILDA_LONG_THREADS=256 ONILDA_LONG_GROUP=8" \
  > gpurun_out/r2_hub1184.log 2>&1
grep -v "^\[onilda\]" gpurun_out/r2_hub1184.log
timeout 900 ncu --set full --import-source on --clock-control none -k regex:estep_kernel --launch-skip 1 -c 1 \
  -o gpurun_out/r2_long_alone python tools/estep_sweep.py --iters 2 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" > gpurun_out/r2_ncu_long.log 2>&1
tail -3 gpurun_out/r2_ncu_long.log
