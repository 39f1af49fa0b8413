mkdir -p gpurun_out
python -m pytest tests/test_move.py tests/test_csv_order.py -m gpu -x -q > gpurun_out/pytest_gpu7.log 2>&1; tail -3 gpurun_out/pytest_gpu7.log
echo "== occ4 cap default"; timeout 900 python profiles/scripts/configs_sweep.py --skip-cfg5 --cfg4-move 2> gpurun_out/exp6_a.err | tee gpurun_out/exp6_a.jsonl | cut -c1-60,200-420
echo "== occ8 cap 12"; EDSBWT_MOVE_OCC8=1 timeout 900 python profiles/scripts/configs_sweep.py --skip-cfg5 --cfg4-move --fused-cap 12 2> gpurun_out/exp6_b.err | tee gpurun_out/exp6_b.jsonl | cut -c1-60,200-420
tail -3 gpurun_out/exp6_a.err
