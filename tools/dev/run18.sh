python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "log_space" --tb=short 2>&1 | tail -25
