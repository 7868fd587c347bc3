L=open-network-insight_b200/lib
for v in base nb2 nopin nb2nopin; do
  cp $L/variants/libonilda_$v.so $L/libonilda.so
  echo "== variant $v"
  timeout 300 python tools/estep_sweep.py --iters 3 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" 2>&1 | grep estep_ms
done
cp $L/variants/libonilda_base.so $L/libonilda.so
timeout 900 ncu --set full --import-source on --clock-control none -k regex:estep_kernel --launch-skip 1 -c 1 \
  -o gpurun_out/r2_long_alone2 python tools/estep_sweep.py --iters 2 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" > gpurun_out/r2_ncu_long2.log 2>&1
tail -2 gpurun_out/r2_ncu_long2.log
