mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_gpu3.log 2>&1; tail -3 gpurun_out/pytest_gpu3.log
for mbp in 250 120 60; do echo "== mbp=$mbp"; python bench.py --no-cpu --steps 5 --mbp $mbp > gpurun_out/exp2_mbp$mbp.json 2> gpurun_out/exp2_mbp$mbp.err; python - <<PY
import json
d=json.load(open("gpurun_out/exp2_mbp$mbp.json")); r=d["roofline"]
print(d["value"]/1e9, d["ms_per_step"], r["kernel_ms_per_step"], d["e2e"]["value"]/1e9, "blocks", r["occ_blocks"], "ns/blk", r["kernel_ms_per_step"]*1e6/r["occ_blocks"], "seed_k", d["config"]["seed_k"], "rows", d["config"]["index_rows"])
PY
done
ncu --set full --clock-control none --import-source on -k regex:fused_search_kernel -s 4 -c 1 -f -o gpurun_out/fused_r01b python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/ncu_full.log 2>&1; tail -2 gpurun_out/ncu_full.log
