mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_final.log 2>&1; tail -3 gpurun_out/pytest_gpu_final.log
python bench.py > gpurun_out/bench_final_n1.json 2> gpurun_out/bench_final_n1.err; cut -c1-200 gpurun_out/bench_final_n1.json
ncu --nvtx --nvtx-include "timed/" --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v8.csv python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/ncu_launches_v8.log 2>&1
