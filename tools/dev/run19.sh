python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python tools/estep_sweep.py --iters 3 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" 2>&1 | grep "estep_ms\|FAILED\|rror"
ONILDA_TIMING=1 timeout 900 python tools/estep_sweep.py --iters 21 "" > gpurun_out/r2_sweep19.log 2>&1
awk '/^\[onilda\]/{last=$0} /estep_ms/{print substr($0,1,260); print "    " last}' gpurun_out/r2_sweep19.log
