mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu11.log 2>&1; tail -3 gpurun_out/pytest_gpu11.log
timeout 600 python profiles/scripts/configs_sweep.py --skip-cfg5 2> gpurun_out/exp11.err | cut -c1-40,140-420
