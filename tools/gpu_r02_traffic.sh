#!/bin/bash
# per-launch ncu counters of the trace kernels for profiles/r02_traffic.json (tied to the source hash), then the bench line
O=gpurun_out; mkdir -p $O
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum
timeout 400 ncu --metrics $M --clock-control none -k regex:'k_trace_(any|closest)' --launch-skip 4 --launch-count 4 --csv --log-file $O/fin_traffic_config3.csv \
    python bench.py --workload config3 --steps 1 --warmup 3 --no-cpu-baseline > $O/fin_traffic_run.log 2>&1; echo "ncu traffic exit $?"
python tools/ncu_traffic.py $O/fin_traffic_config3.csv config3 profiles/r02_traffic.json > /dev/null; echo "traffic json exit $?"
timeout 400 python bench.py > $O/fin_bench_config3.json 2> $O/fin_bench_config3.err; echo "bench config3 exit $?"
python -c "
import json; d=json.load(open('$O/fin_bench_config3.json')); r=d['roofline']
print(round(d['value']), round(d['e2e']['value']), d['frame_fnv'], r['issue_frac'], r['dram_frac'], r['thread_inst_per_shadow_ray'])"
