python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/r2_bench_n1d.json 2> gpurun_out/r2_bench_n1d.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r2_bench_n1d.err
python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/r2_bench_ref_d.json 2> gpurun_out/r2_bench_ref_d.err; echo "ref rc=$?"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_n1d.json'))
print({k:d[k] for k in ('value','ms_per_step','n_gpus','steps','warmup','gpu_launches')}, d['e2e'], d['roofline']['frac'], d['clocks'])
print(d['em']['estep_ms'])
print(d['parity']['within_bars'], d['cpu_baseline'])
r=json.load(open('gpurun_out/r2_bench_ref_d.json')); print({k:r.get(k) for k in ('impl','value','unit','ms_per_step')})
PY
