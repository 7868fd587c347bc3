python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 18 --warmup 3 > gpurun_out/r2_bench_n1c.json 2> gpurun_out/r2_bench_n1c.err; tail -c 600 gpurun_out/r2_bench_n1c.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_n1c.json'))
print({k:d[k] for k in ('value','ms_per_step','n_gpus')}, d['e2e']['value'], d['roofline'])
print(d['em']['estep_ms'])
print(d.get('parity'))
PY
