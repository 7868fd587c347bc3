L=open-network-insight_b200/lib
for v in shflsum smemsum; do
  cp $L/variants/libonilda_$v.so $L/libonilda.so
  echo "== variant $v"
  timeout 300 python tools/estep_sweep.py --iters 3 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" 2>&1 | grep "estep_ms\|FAILED\|rror"
  ONILDA_TIMING=1 timeout 600 python tools/estep_sweep.py --iters 6 "ONILDA_BULK_THREADS=128" "ONILDA_BULK_THREADS=128 ONILDA_LONG_THREADS=128" > gpurun_out/r2_sweep8_$v.log 2>&1
  awk '/^\[onilda\]/{last=$0} /estep_ms/{print substr($0,1,150); print "    " last}' gpurun_out/r2_sweep8_$v.log
done
python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
