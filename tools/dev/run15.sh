python -m pytest tests -m gpu -x -q 2>&1 | tail -4
( time python bench.py > gpurun_out/r2_bench15.json 2> gpurun_out/r2_bench15.err ) 2>&1 | grep real
tail -3 gpurun_out/r2_bench15.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench15.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['value'])
print('parity', json.dumps(d['parity'])[:1500])
print('cpu', json.dumps(d['cpu_baseline'])[:1200])
print('cfg1', d['cfg1'])
print('scoring', {k:d['scoring'][k] for k in ('events_per_sec','score_kernel_ms')}, d['scoring'].get('cpu_baseline'))
PY
( time python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/r2_bench15_ref.json 2> gpurun_out/r2_bench15_ref.err ) 2>&1 | grep real
cat gpurun_out/r2_bench15_ref.json | cut -c1-900
