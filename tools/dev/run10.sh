python -m pytest tests -m gpu -x -q 2>&1 | tail -3
ONILDA_TIMING=1 python bench.py --steps 18 --warmup 3 > gpurun_out/r2_bench10.json 2> gpurun_out/r2_bench10.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench10.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e'])
print(d['em']['estep_ms'])
print(d['roofline']['frac'], d['cpu_baseline'])
PY
grep "E-step" gpurun_out/r2_bench10.err | awk 'NR>3 && NR<=24' | cut -c1-120
