# Round 2, GPU call 22: the whole GPU suite on the final tree
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_final2.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_pytest_gpu_final2.log | cut -c1-300
