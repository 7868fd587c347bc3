L=open-network-insight_b200/lib
for v in mma3 mma1; do
cp $L/variants/libonilda_$v.so $L/libonilda.so
echo "== $v"; python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
timeout 600 python tools/estep_sweep.py --iters 21 "" 2>&1 | grep "estep_ms\|rror\|FAIL"
done
cp $L/variants/libonilda_mma3.so $L/libonilda.so
