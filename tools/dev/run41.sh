L=open-network-insight_b200/lib
cp $L/variants/libonilda_mma1.so $L/libonilda.so
echo "== mma1 (bulk with shuffles)"; timeout 600 python tools/estep_sweep.py --iters 21 "" 2>&1 | grep "estep_ms\|rror\|FAIL"
cp $L/variants/libonilda_mma3.so $L/libonilda.so
echo "== mma3"; timeout 600 python tools/estep_sweep.py --iters 21 "" 2>&1 | grep "estep_ms\|rror\|FAIL"
echo "== 592 hubs alone"; timeout 300 python tools/estep_sweep.py --iters 3 --num-docs 592 --hub-frac 1.0 "" 2>&1 | grep "estep_ms\|FAILED\|rror"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:estep_long -c 1 --launch-skip 1 \
  -o gpurun_out/r2_long_alone5 python tools/estep_sweep.py --iters 2 --num-docs 592 --hub-frac 1.0 "" > gpurun_out/r2_ncu_long5.log 2>&1
tail -2 gpurun_out/r2_ncu_long5.log
