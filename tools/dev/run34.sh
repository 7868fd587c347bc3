python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python tools/estep_sweep.py --iters 8 "" 2>&1 | grep "estep_ms\|rror\|FAIL"
python bench.py --workload proxy-5M-docs-1B-tokens-V500k-K50 --scale 0.2 --steps 6 --warmup 3 --no-scoring --no-grouping --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step']); print(d['em']['estep_ms'])"
ONILDA_LONG_GENERIC=1 python bench.py --workload proxy-5M-docs-1B-tokens-V500k-K50 --scale 0.2 --steps 6 --warmup 3 --no-scoring --no-grouping --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('generic long kernel:', d['value'], d['ms_per_step']); print(d['em']['estep_ms'])"
