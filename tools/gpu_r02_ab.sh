#!/bin/bash
# A/B on ONE box: bench config 3 alternating between library builds (VKRT_B200_LIB), then per-launch ncu counters of each.
# usage: gpu_r02_ab.sh TAG variantA variantB ...   (variant "main" = libvkrt_b200.so, else libvkrt_b200_<variant>.so)
TAG=$1; shift
O=gpurun_out; mkdir -p $O
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum
lib() { if [ "$1" = main ]; then echo $PWD/vkrt_b200/csrc/libvkrt_b200.so; else echo $PWD/vkrt_b200/csrc/libvkrt_b200_$1.so; fi; }
for rep in 1 2 3; do
  for v in "$@"; do
    VKRT_B200_LIB=$(lib $v) timeout 300 python bench.py --no-cpu-baseline --steps 10 > $O/${TAG}_${v}_r${rep}.json 2> $O/${TAG}_${v}_r${rep}.err
    python - <<PY
import json
d=json.load(open("$O/${TAG}_${v}_r${rep}.json")); r=d["roofline"]
print("$v rep $rep", round(d["value"]), "ms", round(d["ms_per_step"],2), "fnv", d["frame_fnv"], "stage", [round(x,2) for x in r["stage_ms_each"][:8]], "n", round(r["n_node_any"],2), round(r["n_tri_any"],2), "clk", d["clocks"]["sm_mhz"])
PY
  done
done
for v in "$@"; do
  VKRT_B200_LIB=$(lib $v) timeout 400 ncu --metrics $M --clock-control none -k regex:'k_trace_(any|closest)' --launch-skip 4 --launch-count 4 --csv --log-file $O/${TAG}_${v}_traffic.csv \
      python bench.py --workload config3 --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_${v}_traffic.log 2>&1; echo "ncu $v exit $?"
done
