timeout 900 ncu --set full --import-source on --clock-control none -k regex:estep_long -c 1 --launch-skip 1 \
  -o gpurun_out/r2_long_alone7 python tools/estep_sweep.py --iters 2 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" > gpurun_out/r2_ncu_long7.log 2>&1
tail -2 gpurun_out/r2_ncu_long7.log
