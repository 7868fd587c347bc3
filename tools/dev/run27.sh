timeout 900 python tools/estep_sweep.py --iters 21 "" "" "" "ONILDA_BULK_PER_SM=2" 2>&1 | grep "estep_ms\|rror\|FAIL"
python bench.py --steps 18 --warmup 3 --no-scoring --no-grouping --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value']); print(d['em']['estep_ms'])"
