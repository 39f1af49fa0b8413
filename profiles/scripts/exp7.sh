mkdir -p gpurun_out
for v in "8 192" "10 192" "12 192" "14 192" "16 192" "10 48" "12 48" "12 96"; do set -- $v; python bench.py --no-cpu --steps 5 --fused-cap $1 --fused-cap2 $2 > gpurun_out/exp7_$1_$2.json 2> gpurun_out/exp7_$1_$2.err; python - <<PY
import json
d=json.load(open("gpurun_out/exp7_$1_$2.json")); r=d["roofline"]
print("cap $1 cap2 $2:", round(d["value"]/1e9,3), "step", round(d["ms_per_step"],3), "fused", round(r["kernel_ms_per_step"],3), "ovf", r["overflow_patterns"], "launches", d["gpu_launches"], "e2e", round(d["e2e"]["value"]/1e9,3))
PY
done
