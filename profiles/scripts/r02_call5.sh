# Round 2, GPU call 5: item kernel with deferred straddles: A/B over the batching period, parity subset, ncu of period 3
mkdir -p gpurun_out
timeout 900 python profiles/scripts/r02_items_ab.py --variants step,items:1,items:2,items:3,items:4 > gpurun_out/r02_items_ab5.jsonl 2> gpurun_out/r02_items_ab5.err; echo "ab rc=$?"; cat gpurun_out/r02_items_ab5.jsonl | cut -c1-330; tail -3 gpurun_out/r02_items_ab5.err
EDSBWT_ITEMS_PERIOD=3 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or random_vs_port or packed or trace or fixed_length" > gpurun_out/r02_pytest_gpu_5.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02_pytest_gpu_5.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:search_items_kernel -s 4 -c 1 -f -o gpurun_out/items_r02_d python profiles/scripts/r02_items_ab.py --variants items:3 --steps 3 > gpurun_out/r02_ncu_items5.log 2>&1; tail -2 gpurun_out/r02_ncu_items5.log
