# Round 2, GPU call 17: wide alphabets through the step-loop kernel (3-bit instantiation) + radix-sigma seed keys; cfg4-scale measurement; 2-bit regression
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_wide_alphabet.py tests/test_zzz_no_alloc.py -m gpu -q > gpurun_out/r02_pytest_wide.log 2>&1; echo "wide rc=$?"; tail -30 gpurun_out/r02_pytest_wide.log | cut -c1-300
timeout 1200 python profiles/scripts/r02_wide_measure.py 250 10000000 0.005 > gpurun_out/r02_wide_alphabet.jsonl 2> gpurun_out/r02_wide_alphabet.err; echo "measure rc=$?"; cat gpurun_out/r02_wide_alphabet.jsonl | cut -c1-1800; tail -5 gpurun_out/r02_wide_alphabet.err | cut -c1-300
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_csv_order.py -m gpu -x -q -k "golden or trace or random_vs_port" > gpurun_out/r02_pytest_regress.log 2>&1; echo "regress rc=$?"; tail -4 gpurun_out/r02_pytest_regress.log
