mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:fused_search_kernel -s 4 -c 1 -f -o gpurun_out/fused_r01c python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/ncu_full_c.log 2>&1; tail -2 gpurun_out/ncu_full_c.log
