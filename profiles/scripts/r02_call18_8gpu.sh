# Round 2, GPU call 18 (8 x B200): the bench as the driver launches it at N = 8 and N = 4 (peer gather inside `value`)
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02_topo_8gpu.txt 2>&1
for N in 8 4; do
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err; echo "bench N=$N rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r02_bench_n$N.json").read().strip().splitlines()[-1])
    g = d.get("gather", {})
    print("N=$N value %.3f G  ms/step %.3f  no-gather %.3f G  e2e %.3f G  parity %s" % (d["value"] / 1e9, d["ms_per_step"], d.get("value_no_gather", 0) / 1e9, d["e2e"]["value"] / 1e9, str(g.get("gather_parity"))[:60]))
except Exception as e:
    print("N=$N no line:", e)
PY
  tail -3 gpurun_out/r02_bench_n$N.err | cut -c1-300
done
