# Round 2, GPU call 14: the no-allocation entry point alone in a fresh process, then the whole GPU suite (it runs last there),
# smoke(), full bench, reference arm, launch list of the timed region
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zzz_no_alloc.py -m gpu -x -q > gpurun_out/r02_pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -15 gpurun_out/r02_pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_pytest_gpu_final.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/r02_smoke_final.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r02_smoke_final.log
timeout 900 python bench.py --steps 20 > gpurun_out/r02_bench_n1_final.json 2> gpurun_out/r02_bench_n1_final.err; echo "bench rc=$?"; cut -c1-1500 gpurun_out/r02_bench_n1_final.json; tail -3 gpurun_out/r02_bench_n1_final.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_reference_arm.err; echo "ref rc=$?"; cut -c1-600 gpurun_out/r02_bench_reference_arm.json
timeout 900 ncu --nvtx --nvtx-include "timed/" --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/r02_launches_run.log 2>&1; echo "ncu rc=$?"; wc -l gpurun_out/r02_launches.csv; tail -3 gpurun_out/r02_launches.csv | cut -c1-300
