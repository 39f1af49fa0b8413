mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu5.log 2>&1; tail -3 gpurun_out/pytest_gpu5.log
for ch in 0 1048576 524288 262144; do echo "== chunk=$ch"; python bench.py --no-cpu --steps 5 --chunk $ch > gpurun_out/exp4_chunk$ch.json 2> gpurun_out/exp4_chunk$ch.err; python - <<PY
import json
d=json.load(open("gpurun_out/exp4_chunk$ch.json")); e=d["e2e"]
print(d["value"]/1e9, d["ms_per_step"], "e2e", e["value"]/1e9, e["ms_per_step"], "with offsets", e["with_offset_array"]["value"]/1e9, e["with_offset_array"]["ms_per_step"])
PY
done
