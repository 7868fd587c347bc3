python -m pytest tests -m gpu -x -q 2>&1 | tail -3
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/r2_bench16_n2.json 2> gpurun_out/r2_bench16_n2.err ) 2>&1 | grep real
grep -i "error\|failed\|Traceback" gpurun_out/r2_bench16_n2.err | head
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench16_n2.json'))
print({k:d[k] for k in ('value','ms_per_step','n_gpus')}, d['e2e']['value'])
print('strong', json.dumps(d['strong'])[:2500])
print('dns', json.dumps(d['strong_dns'])[:600])
print('scoring', json.dumps(d['scoring'])[:600])
print('exchange', d['exchange_in_step'])
PY
