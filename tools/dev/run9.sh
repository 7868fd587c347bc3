L=open-network-insight_b200/lib
for v in v1 v2 v3; do
  cp $L/variants/libonilda_$v.so $L/libonilda.so
  echo "== variant $v"
  ONILDA_TIMING=1 timeout 600 python tools/estep_sweep.py --iters 6 "ONILDA_BULK_THREADS=128" "ONILDA_BULK_THREADS=256" "ONILDA_BULK_THREADS=128 ONILDA_LONG_THREADS=192" "ONILDA_BULK_THREADS=192" > gpurun_out/r2_sweep9_$v.log 2>&1
  awk '/^\[onilda\]/{last=$0} /estep_ms/{print substr($0,1,150); print "    " last}' gpurun_out/r2_sweep9_$v.log
done
