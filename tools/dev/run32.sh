( time python bench.py > gpurun_out/r2_bench_final_1gpu.json 2> gpurun_out/r2_bench_final_1gpu.err ) 2>&1 | grep real
( time python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/r2_bench_final_ref.json 2> gpurun_out/r2_bench_final_ref.err ) 2>&1 | grep real
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_final_1gpu.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['roofline']['frac'], d['clocks'])
print(d['em']['estep_ms'])
print('parity', {k:d['parity'][k] for k in ('iterations_equal','lik_rel_err_max','gamma_max_rel','beta_max_rel','top1000_overlap','within_bars')})
print('cpu', d['cpu_baseline']['value'], d['cpu_baseline']['gpu_same_sample']['value'], d['cpu_baseline']['single_process']['value'])
print('scoring', d['scoring']['events_per_sec'], d['scoring']['score_kernel_ms'], d['scoring']['roofline']['frac'], d['scoring']['e2e']['value'])
print('grouping', d['grouping']['triples_per_sec'], 'cfg1', d['cfg1']['lda_est_wall_s'], d['cfg1']['lik_rel_err_max'])
r=json.load(open('gpurun_out/r2_bench_final_ref.json')); print('ref', r['value'], r['cpu_baseline']['cores'])
PY
