set -x
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench20_short.json 2> gpurun_out/r2_bench20_short.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"estep|column|normalise|doc_partial|doc_final|score|select|radix|hist|scan|totals|offsets|tables_from|export|theta|phi" -c 400 --csv --log-file gpurun_out/launches_r2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_ncu_launches.log 2>&1; echo "ncu1 rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:estep -c 2 --launch-skip 6 -o gpurun_out/prof_r2_estep python tools/estep_sweep.py --iters 5 > gpurun_out/r2_ncu_estep.log 2>&1; echo "ncu2 rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:score_pairs -c 1 --launch-skip 8 -o gpurun_out/prof_r2_score python tools/score_time.py --events 200000000 > gpurun_out/r2_ncu_score.log 2>&1; echo "ncu3 rc=$?"
ls -la gpurun_out/*.ncu-rep
