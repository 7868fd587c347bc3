L=open-network-insight_b200/lib
cp $L/variants/libonilda_base.so $L/libonilda.so
echo "== base"; timeout 600 python tools/estep_sweep.py --iters 21 "" 2>&1 | grep "estep_ms\|rror\|FAIL"
cp $L/variants/libonilda_r384.so $L/libonilda.so
echo "== r384"; timeout 900 python tools/estep_sweep.py --iters 21 "ONILDA_BULK_THREADS=96" "ONILDA_BULK_THREADS=192" "ONILDA_BULK_THREADS=128" "ONILDA_BULK_THREADS=384" 2>&1 | grep "estep_ms\|rror\|FAIL"
cp $L/variants/libonilda_base.so $L/libonilda.so
