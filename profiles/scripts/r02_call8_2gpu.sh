# Round 2, 2-GPU call: the in-process multi-GPU search (NCCL and peer-copy paths), the torchrun gather over NCCL, bench --gpus 2
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02_topo_2gpu.txt 2>&1
timeout 1200 python -m pytest tests/test_gpu_parity.py::test_occurrence_overflow_halves_the_chunk tests/test_multi.py tests/test_sharded.py -m gpu -q > gpurun_out/r02_pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02_pytest_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02_bench_n2.json 2> gpurun_out/r02_bench_n2.err; echo "bench2 rc=$?"; cut -c1-1500 gpurun_out/r02_bench_n2.json; tail -5 gpurun_out/r02_bench_n2.err
timeout 600 python profiles/scripts/r02_multi_cabi.py > gpurun_out/r02_multi_cabi_n2.jsonl 2> gpurun_out/r02_multi_cabi_n2.err; echo "multi rc=$?"; cat gpurun_out/r02_multi_cabi_n2.jsonl; tail -3 gpurun_out/r02_multi_cabi_n2.err
