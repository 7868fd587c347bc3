python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 18 --warmup 3 --no-scoring --no-grouping --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value']); print(d['em']['estep_ms']); print(d['em']['word_steps_per_iter'])"
