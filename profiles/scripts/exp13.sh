mkdir -p gpurun_out
for cap in 0 12 13; do python bench.py --no-cpu --steps 10 --fused-cap $cap > gpurun_out/exp13_$cap.json 2> gpurun_out/exp13_$cap.err; python - <<PY
import json
d=json.load(open("gpurun_out/exp13_$cap.json")); r=d["roofline"]
print("cap=$cap", round(d["value"]/1e9,3), "step", round(d["ms_per_step"],3), "fused", round(r["kernel_ms_per_step"],3), "launches", d["gpu_launches"], "ovf", r["overflow_patterns"])
PY
done
