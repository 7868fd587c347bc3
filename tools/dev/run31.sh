python bench.py --workload proxy-5M-docs-1B-tokens-V500k-K50 --scale 0.2 --steps 6 --warmup 3 --no-scoring --no-grouping --no-cpu-baseline > gpurun_out/r2_bench_proxy02.json 2> gpurun_out/r2_bench_proxy02.err; echo rc=$?
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_proxy02.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['value'], d['config']['nnz'], d['config']['tokens'], d['roofline']['frac'])
print(d['em']['estep_ms'])
PY
tail -2 gpurun_out/r2_bench_proxy02.err
