# Round 2, GPU call 20: run_search with one host round trip per chunk and a deferred tail: host-buffer parity tests + e2e chunk sweep
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_csv_order.py tests/test_wide_alphabet.py -m gpu -x -q > gpurun_out/r02_pytest_async.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02_pytest_async.log | cut -c1-300
timeout 600 python profiles/scripts/r02_e2e_chunks.py 524288,1048576,2097152,3145728,4194304,10000000 > gpurun_out/r02_e2e_chunks2.jsonl 2>gpurun_out/r02_e2e_chunks2.err; cat gpurun_out/r02_e2e_chunks2.jsonl | cut -c1-330; tail -3 gpurun_out/r02_e2e_chunks2.err
