L=open-network-insight_b200/lib
for v in pairs0 pairs1 pairs0 pairs1; do
cp $L/variants/libonilda_$v.so $L/libonilda.so
echo "== $v"; timeout 600 python tools/estep_sweep.py --iters 21 "" "" 2>&1 | grep "estep_ms\|rror\|FAIL" | tail -1
done
