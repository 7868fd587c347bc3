python -m pytest tests/test_gpu_parity.py tests/test_device_math.py -m gpu -x -q 2>&1 | tail -2
timeout 1200 python tools/estep_sweep.py --iters 21 "" "" 2>&1 | grep "estep_ms\|rror\|FAIL"
echo "== 592 hubs alone"; timeout 300 python tools/estep_sweep.py --iters 3 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" 2>&1 | grep "estep_ms\|FAILED\|rror"
