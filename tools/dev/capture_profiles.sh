set -x
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench59_short.json 2> gpurun_out/r2_bench59_short.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"estep|column|normalise|doc_partial|doc_final|score|select|radix|hist|scan|totals|offsets|tables_from|export|theta|phi" -c 400 --csv --log-file gpurun_out/launches_r2b.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_ncu_launches_b.log 2>&1; echo "ncu1 rc=$?"
python tools/estep_sweep.py --iters 5 > gpurun_out/r2_sweep59.log 2>&1; echo "sweep rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:estep_long -c 1 --launch-skip 3 -o gpurun_out/prof_r2b_estep python tools/estep_sweep.py --iters 5 > gpurun_out/r2_ncu_estep_b.log 2>&1; echo "ncu2 rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:estep_long -c 1 --launch-skip 1 -o gpurun_out/r2_long_alone8 python tools/estep_sweep.py --iters 2 --num-docs 592 --hub-frac 1.0 "ONILDA_LONG_THREADS=128" > gpurun_out/r2_ncu_long8.log 2>&1; echo "ncu3 rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:estep_kernel -c 1 --launch-skip 3 -o gpurun_out/prof_r2b_estep_bulk python tools/estep_sweep.py --iters 5 > gpurun_out/r2_ncu_estep_bulk_b.log 2>&1; echo "ncu4 rc=$?"
ls -la gpurun_out/*.ncu-rep | tail -4
