# Round 2, GPU call 2: item kernel v2 (seeds resolved at pool grab, finish + refill merged, one code path for both passes)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_3.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02_pytest_gpu_3.log
timeout 900 python profiles/scripts/r02_items_ab.py --variants step,items:10,items:9,items:8 > gpurun_out/r02_items_ab3.jsonl 2> gpurun_out/r02_items_ab3.err; echo "ab rc=$?"; cat gpurun_out/r02_items_ab3.jsonl; tail -3 gpurun_out/r02_items_ab3.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:search_items_kernel -s 4 -c 1 -f -o gpurun_out/items_r02_c python profiles/scripts/r02_items_ab.py --variants items:10 --steps 3 > gpurun_out/r02_ncu_items3.log 2>&1; tail -2 gpurun_out/r02_ncu_items2.log
