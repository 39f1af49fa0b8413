# Round 2, GPU call 16: wide-alphabet CLI test + the wide path at cfg4 scale (N in 0.5 % of the letters)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_wide_alphabet.py -m gpu -q -k "cli" > gpurun_out/r02_pytest_wide_cli.log 2>&1; echo "wide cli rc=$?"; tail -30 gpurun_out/r02_pytest_wide_cli.log | cut -c1-300
timeout 1200 python profiles/scripts/r02_wide_measure.py 250 2000000 0.005 > gpurun_out/r02_wide_alphabet.jsonl 2> gpurun_out/r02_wide_alphabet.err; echo "measure rc=$?"; cat gpurun_out/r02_wide_alphabet.jsonl | cut -c1-1500; tail -5 gpurun_out/r02_wide_alphabet.err
