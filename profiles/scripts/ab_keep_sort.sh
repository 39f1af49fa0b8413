mkdir -p gpurun_out
python -m pytest tests/test_csv_order.py tests/test_gpu_parity.py tests/test_move.py -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; tail -15 gpurun_out/pytest_gpu2.log
for v in "0 0" "1 0" "0 1" "1 1"; do set -- $v; echo "== KEEP=$1 SORT=$2"; EDSBWT_OCC_KEEP=$1 EDSBWT_SORT=$2 python bench.py --no-cpu --steps 5 > gpurun_out/ab_keep$1_sort$2.json 2> gpurun_out/ab_keep$1_sort$2.err; python - <<PY
import json
d=json.load(open("gpurun_out/ab_keep$1_sort$2.json"))
print(d["value"]/1e9, d["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["e2e"]["value"]/1e9)
PY
done
