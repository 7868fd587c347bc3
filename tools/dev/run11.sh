L=open-network-insight_b200/lib
for v in ring16 ring8; do
  cp $L/variants/libonilda_$v.so $L/libonilda.so
  echo "== variant $v"
  timeout 900 python tools/estep_sweep.py --iters 21 "ONILDA_BULK_THREADS=128" "ONILDA_BULK_THREADS=256" 2>&1 | grep estep_ms
done
