mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_csv_order.py -m gpu -x -q > gpurun_out/pytest_gpu15.log 2>&1; tail -3 gpurun_out/pytest_gpu15.log
python bench.py --no-cpu --steps 10 > gpurun_out/exp15.json 2> gpurun_out/exp15.err; python - <<PY
import json
d=json.load(open("gpurun_out/exp15.json")); r=d["roofline"]
print(round(d["value"]/1e9,3), "step", round(d["ms_per_step"],3), "fused", round(r["kernel_ms_per_step"],3), "locate", round(r["locate_kernel_ms_per_step"],3), "launches", d["gpu_launches"])
PY
timeout 300 python profiles/scripts/configs_sweep.py --skip-cfg3 --lens 8,16 2> gpurun_out/exp15s.err | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); c=d['count_only']; o=d['locate']
    print(d['pattern_len'], 'locate', o.get('patterns'), round(o.get('patterns_per_s',0)/1e9,4), 'G/s', round(o.get('ms_per_step',0),3), 'ms, locate kernel', round(o.get('locate_kernel_ms',0),3), 'occ', o.get('occurrences'))
"
