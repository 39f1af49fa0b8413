# Round 2, GPU call 15: alphabets of 5..7 letters (3-bit Occ layout, level-synchronous engine) + regression of the 2-bit goldens
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_wide_alphabet.py -m gpu -q > gpurun_out/r02_pytest_wide.log 2>&1; echo "wide rc=$?"; tail -40 gpurun_out/r02_pytest_wide.log | cut -c1-300
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or trace or random_vs_port" > gpurun_out/r02_pytest_regress.log 2>&1; echo "regress rc=$?"; tail -4 gpurun_out/r02_pytest_regress.log
