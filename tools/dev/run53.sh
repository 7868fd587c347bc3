L=open-network-insight_b200/lib
cp $L/variants/libonilda_pairs1.so $L/libonilda.so
timeout 1500 python tools/estep_sweep.py --iters 21 "" "ONILDA_LONG_THREADS=128" "ONILDA_LONG_THREADS=128 ONILDA_BULK_THREADS=128" "ONILDA_LONG_THREADS=128 ONILDA_BULK_THREADS=384" "ONILDA_LONG_THREADS=192" "ONILDA_LONG_THREADS=128 ONILDA_LONG_LEN=2048" 2>&1 | grep "estep_ms\|rror\|FAIL"
