mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_final.log 2>&1; tail -3 gpurun_out/pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/bench_final_n1.json 2> gpurun_out/bench_final_n1.err; tail -c 200 gpurun_out/bench_final_n1.err; cut -c1-200 gpurun_out/bench_final_n1.json
timeout 900 python profiles/scripts/configs_sweep.py --cfg4-move > gpurun_out/configs_sweep.jsonl 2> gpurun_out/configs_sweep.err; wc -l gpurun_out/configs_sweep.jsonl
ncu --nvtx --nvtx-include "timed/" --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v7.csv python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/ncu_launches_v7.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fused_search_kernel -s 4 -c 1 -f -o gpurun_out/fused_r01g python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/ncu_full_g.log 2>&1; tail -2 gpurun_out/ncu_full_g.log
