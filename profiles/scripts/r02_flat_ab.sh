# Round 2, first GPU call: the experimental warp-flat first pass (fused.cu, FLAT = true) against the shipped kernel.
#   gpurun --timeout 900 -- 'bash profiles/scripts/r02_flat_ab.sh'
# 1. parity of the three variants incl. their counting instantiations (only variant 1, non-counting, ran in round 1), 2. A/B of the bench at cfg4, 3. one ncu capture of the flat kernel.
mkdir -p gpurun_out
EDSBWT_TEST_FLAT=1 python -m pytest tests/test_zz_flat_experimental.py -m gpu -x -q > gpurun_out/r02_flat_parity.log 2>&1; tail -3 gpurun_out/r02_flat_parity.log
for v in 0 1 2 3; do    # 0 shipped step loop, 1 warp-flat at 10 CTAs/SM (48 registers), 2 warp-flat at 12 CTAs/SM (40 registers, 120 B of spills), 3 = 1 + warp-wide link gather
  EDSBWT_FUSED_FLAT=$v python bench.py --no-cpu --steps 20 > gpurun_out/r02_bench_flat$v.json 2> gpurun_out/r02_bench_flat$v.err
  python - <<PY
import json
d = json.load(open("gpurun_out/r02_bench_flat$v.json"))
print("flat=$v  %.3f G patterns/s  step %.3f ms  fused kernel %.3f ms  overflow %d" % (d["value"] / 1e9, d["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["roofline"]["overflow_patterns"]))
PY
done
EDSBWT_FUSED_FLAT=1 ncu --set full --clock-control none --import-source on -k regex:fused_search_kernel -s 4 -c 1 -f -o gpurun_out/fused_r02_flat python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/r02_ncu_flat.log 2>&1; tail -2 gpurun_out/r02_ncu_flat.log
