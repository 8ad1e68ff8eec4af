#!/bin/bash
# Round 2, GPU call B: A/B of a trace-kernel variant: GPU parity tier, bench config 3 (+ others), per-launch ncu counters.
TAG=${1:-b}
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1; echo "gpu tests exit $?" | tee -a $O/${TAG}_gpu_tests.log
tail -5 $O/${TAG}_gpu_tests.log
timeout 400 python bench.py --no-cpu-baseline > $O/${TAG}_bench_config3.json 2> $O/${TAG}_bench_config3.err; echo "bench config3 exit $?"
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench_config3.json")); r=d["roofline"]
print("config3", round(d["value"]), "ms", round(d["ms_per_step"],2), "fnv", d["frame_fnv"], "stage", [round(x,2) for x in r["stage_ms_each"]], "n", round(r["n_node_any"],2), round(r["n_tri_any"],2), round(r["n_node_closest"],2), round(r["n_tri_closest"],2))
PY
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum
timeout 400 ncu --metrics $M --clock-control none -k regex:'k_trace_(any|closest)' --launch-skip 4 --launch-count 4 --csv --log-file $O/${TAG}_traffic_config3.csv \
    python bench.py --workload config3 --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_traffic_run.log 2>&1; echo "ncu traffic exit $?"
for w in config1 config2 config4; do
  timeout 300 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_$w.json 2> $O/${TAG}_bench_$w.err; echo "bench $w exit $?"
done
timeout 400 python bench.py --workload config5 --steps 5 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_config5.json 2> $O/${TAG}_bench_config5.err; echo "bench config5 exit $?"
python - <<PY
import json
for w in ["config1","config2","config4","config5"]:
    try:
        d=json.load(open("$O/${TAG}_bench_%s.json" % w)); r=d["roofline"]
        print(w, round(d["value"]), "ms", round(d["ms_per_step"],2), "n", round(r["n_node_any"],2), round(r["n_tri_any"],2), round(r["n_node_closest"],2), round(r["n_tri_closest"],2))
    except Exception as e:
        print(w, "failed", e)
PY
