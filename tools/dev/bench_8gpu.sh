( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29563 bench.py --gpus 8 --steps 18 --warmup 3 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err ) 2>&1 | grep real
grep -i "error\|failed\|Traceback" gpurun_out/r2_bench_n8.err | head
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_n8.json'))
print({k:d[k] for k in ('value','ms_per_step','n_gpus')}, d['e2e']['value'])
print(d['em']['estep_ms'])
for k in ('strong','strong_dns','strong_proxy'):
    s=d[k]; print(k, {x:s.get(x) for x in ('num_docs','tokens','ms_per_iter','estep_ms_max_over_ranks','tokens_per_sec')}, s.get('exchange_alone',{}).get('reduce_scatter_ms'), s.get('exchange_alone',{}).get('all_gather_x3_ms'), s.get('exchange_alone',{}).get('allreduce_equivalent_bus_gbs'), (s.get('longest_document') or {}).get('lone_warp_estep_ms'))
print('scoring', d['scoring']['sharded']['events_per_sec'], d['scoring']['sharded']['ms'])
PY
