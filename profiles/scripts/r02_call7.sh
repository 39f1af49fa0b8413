mkdir -p gpurun_out
timeout 900 python profiles/scripts/r02_items_ab.py --variants step,items,items-pack > gpurun_out/r02_items_ab7.jsonl 2> gpurun_out/r02_items_ab7.err; echo "ab rc=$?"; cat gpurun_out/r02_items_ab7.jsonl | cut -c1-230; tail -3 gpurun_out/r02_items_ab7.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:search_items_kernel -s 4 -c 1 -f -o gpurun_out/items_r02_f python profiles/scripts/r02_items_ab.py --variants items --steps 3 > gpurun_out/r02_ncu_items7.log 2>&1; tail -2 gpurun_out/r02_ncu_items7.log
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02_pytest_gpu_7.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_7.log
