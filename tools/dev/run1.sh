set -x
nvidia-smi --query-gpu=name,memory.total --format=csv
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_tests1.log
ONILDA_TIMING=1 timeout 900 python tools/estep_sweep.py --iters 8 \
  "" \
  "ONILDA_BULK_THREADS=384" \
  "ONILDA_BULK_THREADS=128" \
  "ONILDA_LONG_THREADS=128" \
  "ONILDA_LONG_GROUP=8" \
  "ONILDA_LONG_GROUP=8 ONILDA_LONG_THREADS=128" \
  "ONILDA_LONG_GROUP=16" \
  "ONILDA_LONG_LEN=2048" \
  "ONILDA_LONG_LEN=512" \
  "ONILDA_LONG_PER_SM=2 ONILDA_BULK_THREADS=128" \
  > gpurun_out/r2_sweep1.log 2>&1
tail -5 gpurun_out/r2_tests1.log
grep -v "^\[onilda\]" gpurun_out/r2_sweep1.log
