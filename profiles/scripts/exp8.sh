mkdir -p gpurun_out
for v in "0 12" "1 12" "1 10"; do set -- $v; EDSBWT_FUSED_OCC12=$1 python bench.py --no-cpu --steps 5 --fused-cap $2 > gpurun_out/exp8_$1_$2.json 2> gpurun_out/exp8_$1_$2.err; python - <<PY
import json
d=json.load(open("gpurun_out/exp8_$1_$2.json")); r=d["roofline"]
print("occ12=$1 cap $2:", round(d["value"]/1e9,3), "step", round(d["ms_per_step"],3), "fused", round(r["kernel_ms_per_step"],3), "ovf", r["overflow_patterns"], "launches", d["gpu_launches"])
PY
done
ncu --nvtx --nvtx-include "timed/" --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v4.csv python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/ncu_launches_v4.log 2>&1; grep -c fused gpurun_out/launches_v4.csv
