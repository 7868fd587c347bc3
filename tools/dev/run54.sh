L=open-network-insight_b200/lib
for v in pairs0 pairs1; do
cp $L/variants/libonilda_$v.so $L/libonilda.so
echo "== $v"; timeout 600 python tools/estep_sweep.py --iters 21 --hub-frac 0.0 "" "" 2>&1 | grep "estep_ms\|rror\|FAIL" | tail -1
timeout 600 python tools/estep_sweep.py --iters 8 --workload dns-2M-docs-200M-tokens-V50k-K20 --scale 0.5 "" "" 2>&1 | grep "estep_ms\|rror\|FAIL" | tail -1
done
cp $L/variants/libonilda_pairs1.so $L/libonilda.so
