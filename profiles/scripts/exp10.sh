mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_csv_order.py -m gpu -x -q > gpurun_out/pytest_gpu10.log 2>&1; tail -4 gpurun_out/pytest_gpu10.log
timeout 900 python profiles/scripts/configs_sweep.py --skip-cfg3 --lens 8,12,14,16,32 2> gpurun_out/exp10.err | tee gpurun_out/exp10.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); c=d['count_only']; o=d['locate']
    print(d['pattern_len'], 'count-only', round(c['patterns_per_s']/1e9,3), 'G/s', round(c['ms_per_step'],3),'ms ovf',c['overflow_patterns'], '| locate', o.get('patterns'), round(o.get('patterns_per_s',0)/1e9,4), 'G/s', round(o.get('ms_per_step',0),3), 'ms occ', o.get('occurrences'), 'ovf', o.get('overflow_patterns'))
"
tail -3 gpurun_out/exp10.err
