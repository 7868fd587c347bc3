python -m pytest tests -m gpu -x -q -k "two_gpu or second_device or no_alloc" --tb=short 2>&1 | tail -30
