#!/bin/bash
# ncu --set full of one depth-0 launch of k_trace_any and k_trace_closest per library variant (same box).
TAG=$1; shift
O=gpurun_out; mkdir -p $O
lib() { if [ "$1" = main ]; then echo $PWD/vkrt_b200/csrc/libvkrt_b200.so; else echo $PWD/vkrt_b200/csrc/libvkrt_b200_$1.so; fi; }
for v in "$@"; do
  VKRT_B200_LIB=$(lib $v) timeout 600 ncu --set full --import-source on --clock-control none -k regex:'k_trace_(any|closest)' --launch-skip 4 --launch-count 2 \
      -o $O/${TAG}_${v}_full -f python bench.py --workload config3 --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_${v}_full.log 2>&1; echo "ncu full $v exit $?"
done
ls -la $O/*.ncu-rep
