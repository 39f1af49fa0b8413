# Round 2, GPU call 4: batching period A/B of the item kernel, the new tests (occurrence overflow, cache, multi on one GPU, scale), gather sweep
mkdir -p gpurun_out
timeout 900 python profiles/scripts/r02_items_ab.py --variants step,items:1,items:2,items:3,items:4 > gpurun_out/r02_items_ab4.jsonl 2> gpurun_out/r02_items_ab4.err; echo "ab rc=$?"; cat gpurun_out/r02_items_ab4.jsonl; tail -3 gpurun_out/r02_items_ab4.err
timeout 1800 python -m pytest tests/test_gpu_parity.py::test_occurrence_overflow_halves_the_chunk tests/test_cache.py tests/test_multi.py tests/test_zz_configs_scale.py tests/test_move.py tests/test_sharded.py tests/test_csv_order.py tests/test_tools.py tests/test_zz_threads.py -m gpu -q > gpurun_out/r02_pytest_gpu_4.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r02_pytest_gpu_4.log
timeout 900 python profiles/gather_sweep.py > gpurun_out/r02_gather_sweep.txt 2> gpurun_out/r02_gather_sweep.err; echo "gather rc=$?"; tail -12 gpurun_out/r02_gather_sweep.txt
