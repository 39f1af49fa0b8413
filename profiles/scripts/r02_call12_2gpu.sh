mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sharded.py -m gpu -q > gpurun_out/r02_pytest_2gpu_c.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_2gpu_c.log
for cfg in "peer 4" "peer 3" "peer 6"; do
  set -- $cfg
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --gather $1 --gather-chunks $2 > gpurun_out/r02_bench_n2_$1_$2.json 2> gpurun_out/r02_bench_n2_$1_$2.err; echo "bench $cfg rc=$?"
  python - <<PY
import json
d = json.load(open("gpurun_out/r02_bench_n2_$1_$2.json"))
print("$cfg", "value %.3f G  ms/step %.3f  no-gather %.3f G  e2e %.3f G" % (d["value"]/1e9, d["ms_per_step"], d["value_no_gather"]/1e9, d["e2e"]["value"]/1e9), d["gather"]["gather_parity"][:60])
PY
done
