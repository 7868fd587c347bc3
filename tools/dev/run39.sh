nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/wsl tools/warp_sum_latency.cu && /tmp/wsl | tee gpurun_out/warp_sum_latency.txt
