L=open-network-insight_b200/lib
cp $L/variants/libonilda_cur.so $L/libonilda.so
echo "== lean"; timeout 600 python tools/estep_sweep.py --iters 21 --hub-frac 0.0 "" "ONILDA_BULK_THREADS=512" 2>&1 | grep "estep_ms\|rror\|FAIL"
cp $L/variants/libonilda_fat.so $L/libonilda.so
echo "== fat"; timeout 600 python tools/estep_sweep.py --iters 21 --hub-frac 0.0 "ONILDA_BULK_THREADS=384" "ONILDA_BULK_THREADS=192" "ONILDA_BULK_THREADS=128" 2>&1 | grep "estep_ms\|rror\|FAIL"
cp $L/variants/libonilda_cur.so $L/libonilda.so
