mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_csv_order.py -m gpu -x -q > gpurun_out/pytest_gpu4.log 2>&1; tail -3 gpurun_out/pytest_gpu4.log
python bench.py --no-cpu --steps 5 > gpurun_out/exp3.json 2> gpurun_out/exp3.err; python - <<PY
import json
d=json.load(open("gpurun_out/exp3.json")); r=d["roofline"]
print(d["value"]/1e9, d["ms_per_step"], r["kernel_ms_per_step"], d["e2e"]["value"]/1e9, "frac", r["frac"], "ovf", r["overflow_patterns"])
PY
