mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_gpu14.log 2>&1; tail -3 gpurun_out/pytest_gpu14.log
python bench.py --no-cpu --steps 10 > gpurun_out/exp14.json 2> gpurun_out/exp14.err; grep "index open" gpurun_out/exp14.err; python - <<PY
import json
d=json.load(open("gpurun_out/exp14.json")); r=d["roofline"]
print(round(d["value"]/1e9,3), "step", round(d["ms_per_step"],3), "fused", round(r["kernel_ms_per_step"],3), "launches", d["gpu_launches"], "ovf", r["overflow_patterns"])
PY
