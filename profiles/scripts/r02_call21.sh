# Round 2, GPU call 21: four result slots in the chunk pipeline: pipeline tests + e2e chunk sweep
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_zz_threads.py tests/test_wide_alphabet.py -m gpu -x -q -k "search_into or overflow or concurrent or wide_random" > gpurun_out/r02_pytest_pipe.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_pytest_pipe.log | cut -c1-300
timeout 600 python profiles/scripts/r02_e2e_chunks.py 1048576,2097152,3145728,4194304,5242880 > gpurun_out/r02_e2e_chunks3.jsonl 2>gpurun_out/r02_e2e_chunks3.err; cat gpurun_out/r02_e2e_chunks3.jsonl | cut -c1-200; tail -3 gpurun_out/r02_e2e_chunks3.err
