nvidia-smi -L
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "two_gpu or sharded or second_device" 2>&1 | tail -15
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
