python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
timeout 1200 python tools/estep_sweep.py --iters 21 "" 2>&1 | grep "estep_ms\|rror\|FAIL"
