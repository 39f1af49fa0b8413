# Round 2, GPU call 1: parity of the item kernel (whole GPU suite), A/B against the step-loop kernel, one ncu capture.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_1.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02_pytest_gpu_1.log
timeout 900 python profiles/scripts/r02_items_ab.py > gpurun_out/r02_items_ab.jsonl 2> gpurun_out/r02_items_ab.err; echo "ab rc=$?"; cat gpurun_out/r02_items_ab.jsonl; tail -3 gpurun_out/r02_items_ab.err
timeout 600 python bench.py --no-cpu --steps 10 > gpurun_out/r02_bench_items_1.json 2> gpurun_out/r02_bench_items_1.err; echo "bench rc=$?"; cat gpurun_out/r02_bench_items_1.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:search_items_kernel -s 4 -c 1 -f -o gpurun_out/items_r02_a python profiles/scripts/r02_items_ab.py --variants items:10 --steps 3 > gpurun_out/r02_ncu_items.log 2>&1; tail -2 gpurun_out/r02_ncu_items.log
