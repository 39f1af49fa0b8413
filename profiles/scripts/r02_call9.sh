# Round 2, GPU call 9: whole GPU suite on the final kernels + full bench + launch list + ncu captures of both first-pass kernels
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_9.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02_pytest_gpu_9.log
timeout 900 python bench.py --steps 20 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc=$?"; cut -c1-3000 gpurun_out/r02_bench_n1.json; tail -4 gpurun_out/r02_bench_n1.err
