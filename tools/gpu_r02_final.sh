#!/bin/bash
# Round 2 evidence run on ONE B200: GPU tier, bench lines of every config, the ncu captures that profiles/r02_* summarise.
O=gpurun_out; mkdir -p $O
T=fin
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > $O/${T}_gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > $O/${T}_gpu_tests.log 2>&1; echo "gpu tests exit $?"; tail -3 $O/${T}_gpu_tests.log
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum
timeout 400 ncu --metrics $M --clock-control none -k regex:'k_trace_(any|closest)' --launch-skip 4 --launch-count 4 --csv --log-file $O/${T}_traffic_config3.csv \
    python bench.py --workload config3 --steps 1 --warmup 3 --no-cpu-baseline > $O/${T}_traffic_run.log 2>&1; echo "ncu traffic exit $?"
python tools/ncu_traffic.py $O/${T}_traffic_config3.csv config3 profiles/r02_traffic.json > /dev/null; echo "traffic json exit $?"
timeout 400 python bench.py > $O/${T}_bench_config3.json 2> $O/${T}_bench_config3.err; echo "bench config3 exit $?"
for w in config1 config2 config4; do
  timeout 300 python bench.py --workload $w --steps 10 --warmup 3 > $O/${T}_bench_$w.json 2> $O/${T}_bench_$w.err; echo "bench $w exit $?"
done
timeout 400 python bench.py --workload config5 --steps 5 --warmup 3 --no-cpu-baseline > $O/${T}_bench_config5.json 2> $O/${T}_bench_config5.err; echo "bench config5 exit $?"
timeout 300 python bench.py --workload raster --steps 20 --warmup 3 > $O/${T}_bench_raster.json 2> $O/${T}_bench_raster.err; echo "bench raster exit $?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_reference.json 2> $O/${T}_bench_reference.err; echo "reference arm exit $?"
python - <<PY
import json
for w in ["config3","config1","config2","config4","config5"]:
    try:
        d=json.load(open("$O/${T}_bench_%s.json" % w)); r=d["roofline"]
        print(w, round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"],2), "fnv", d["frame_fnv"], "frac", round(r["frac"],3), "issue", r["issue_frac"], "dram", r["dram_frac"], "inst/ray", r["thread_inst_per_shadow_ray"], "build", d["bvh_build"]["warm_ms"])
    except Exception as e:
        print(w, "failed", e)
PY
# launch list of the timed frames only (builder and mip kernels filtered out), then one full capture of the top kernel
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_(raygen|trace|shade|resolve|reset|untile)' --launch-skip 24 --launch-count 48 --csv --log-file $O/${T}_launches_config3.csv \
    python bench.py --workload config3 --steps 2 --warmup 3 --no-cpu-baseline > $O/${T}_launches_run.log 2>&1; echo "ncu launches exit $?"
timeout 600 ncu --set full --import-source on --clock-control none -k regex:'k_trace_(any|closest)' --launch-skip 4 --launch-count 2 -o $O/${T}_trace_full -f \
    python bench.py --workload config3 --steps 1 --warmup 3 --no-cpu-baseline > $O/${T}_trace_full.log 2>&1; echo "ncu full exit $?"
timeout 300 python tools/build_bench.py --flags 0 --sizes 262144,2000000,10000000 > $O/${T}_build_ploc.jsonl 2>&1; echo "build ploc exit $?"
timeout 300 python tools/build_bench.py --flags 1 > $O/${T}_build_lbvh.jsonl 2>&1; echo "build lbvh exit $?"
timeout 300 python tools/build_bench.py --flags 8 --sizes 262144,2000000 > $O/${T}_build_fast_trace.jsonl 2>&1; echo "build fast trace exit $?"
for size in 1 8; do timeout 300 python tools/rank_stage_times.py --size $size 2>&1 | tail -1; done > $O/${T}_rank_stage_times.txt; echo "rank stage times exit $?"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke exit $?"
cp profiles/r02_traffic.json $O/${T}_r02_traffic.json
ls $O | grep "^${T}_" | wc -l
