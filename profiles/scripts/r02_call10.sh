# Round 2, GPU call 10: ncu of the step-loop kernel + launch list of the bench, cfg5 length sweep, MOVE on repetitive EDSs,
# open / cache timings, file -> csv through the executable
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fused_search_kernel -s 4 -c 1 -f -o gpurun_out/fused_r02 python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/r02_ncu_fused.log 2>&1; tail -1 gpurun_out/r02_ncu_fused.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/r02_ncu_launches.log 2>&1; tail -1 gpurun_out/r02_ncu_launches.log
for m in 8 16 32 64 128 256; do
  timeout 900 python bench.py --config cfg5 --len $m --steps 5 --total ${CFG5_TOTAL:-100000000} >> gpurun_out/r02_cfg5_sweep.jsonl 2> gpurun_out/r02_cfg5_m$m.err; echo "cfg5 m=$m rc=$?"; tail -1 gpurun_out/r02_cfg5_sweep.jsonl | cut -c1-900
done
timeout 900 python profiles/scripts/r02_move_repetitive.py > gpurun_out/r02_move_repetitive.jsonl 2> gpurun_out/r02_move_repetitive.err; echo "move rc=$?"; cat gpurun_out/r02_move_repetitive.jsonl | cut -c1-700
timeout 600 python profiles/scripts/r02_open_timing.py > gpurun_out/r02_open_timing.json 2> gpurun_out/r02_open_timing.err; echo "open rc=$?"; cat gpurun_out/r02_open_timing.json
timeout 900 python profiles/scripts/cli_cfg4.py > gpurun_out/r02_cli_cfg4.json 2> gpurun_out/r02_cli_cfg4.err; echo "cli rc=$?"; cut -c1-1500 gpurun_out/r02_cli_cfg4.json
