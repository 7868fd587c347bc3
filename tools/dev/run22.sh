( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 18 --warmup 3 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err ) 2>&1 | grep real
grep -i "error\|failed\|Traceback" gpurun_out/r2_bench_n8.err | head
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_n8.json'))
print({k:d[k] for k in ('value','ms_per_step','n_gpus')}, d['e2e']['value'])
s=d['strong']; print('strong', {k:s.get(k) for k in ('ms_per_iter','estep_ms_max_over_ranks','tokens_per_sec','roofline_frac')}, s.get('exchange_alone'), s.get('longest_document'))
s=d['strong_dns']; print('dns', {k:s.get(k) for k in ('ms_per_iter','tokens_per_sec','roofline_frac')} if s else None)
print('scoring', json.dumps(d['scoring'])[:400])
print('exchange', d['exchange_in_step'])
PY
