L=open-network-insight_b200/lib
for v in base A B; do
  cp $L/variants/libonilda_$v.so $L/libonilda.so
  echo "== variant $v"
  python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "single_estep or sweep_counts or full_em or cfg1" 2>&1 | tail -1
  timeout 600 python tools/estep_sweep.py --iters 21 "" 2>&1 | grep "estep_ms\|rror\|FAIL"
done
cp $L/variants/libonilda_base.so $L/libonilda.so
