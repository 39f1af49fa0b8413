mkdir -p gpurun_out
TAG=${1:-x}
timeout 900 python profiles/scripts/r02_items_ab.py --variants step,items-pack --steps 10 > gpurun_out/r02_items_ab_$TAG.jsonl 2> gpurun_out/r02_items_ab_$TAG.err; echo "ab rc=$?"; cat gpurun_out/r02_items_ab_$TAG.jsonl | cut -c1-230; tail -3 gpurun_out/r02_items_ab_$TAG.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:search_items_kernel -s 4 -c 1 -f -o gpurun_out/items_r02_$TAG python profiles/scripts/r02_items_ab.py --variants items-pack --steps 3 > gpurun_out/r02_ncu_items_$TAG.log 2>&1; tail -1 gpurun_out/r02_ncu_items_$TAG.log
