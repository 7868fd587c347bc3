timeout 900 ncu --set full --import-source on --clock-control none -k regex:score_pairs -c 1 --launch-skip 3 -o gpurun_out/prof_r2_score python tools/score_time.py --events 200000000 > gpurun_out/r2_ncu_score.log 2>&1; echo "ncu3 rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:estep_kernel -c 1 --launch-skip 3 -o gpurun_out/prof_r2_estep_bulk python tools/estep_sweep.py --iters 5 > gpurun_out/r2_ncu_estep_bulk.log 2>&1; echo "ncu4 rc=$?"
ls -la gpurun_out/*.ncu-rep
