mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_csv_order.py -m gpu -x -q > gpurun_out/pytest_gpu9.log 2>&1; tail -3 gpurun_out/pytest_gpu9.log
python bench.py --no-cpu --steps 5 > gpurun_out/exp9.json 2> gpurun_out/exp9.err; python - <<PY
import json
d=json.load(open("gpurun_out/exp9.json")); r=d["roofline"]
print(round(d["value"]/1e9,3), "step", round(d["ms_per_step"],3), "fused", round(r["kernel_ms_per_step"],3), "locate", round(r["locate_kernel_ms_per_step"],3), "e2e", round(d["e2e"]["value"]/1e9,3), "bytes", d["config"]["l2"])
PY
