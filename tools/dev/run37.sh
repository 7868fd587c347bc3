python tools/bundle_waste.py 21 > gpurun_out/bundle_waste.txt 2>&1; tail -25 gpurun_out/bundle_waste.txt
