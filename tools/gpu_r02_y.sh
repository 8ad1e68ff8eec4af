#!/bin/bash
# collapse: how large a level the one-block sweep should take (Sponza-sized builds are launch- and latency-bound)
O=gpurun_out; mkdir -p $O
for nm in 1024 256 128 64 32; do
  for wp in 8 12; do
    for fl in 1 0; do
      echo -n "narrowMax $nm pairs $wp flags $fl: "
      VKRT_BUILD_NARROW_MAX=$nm VKRT_BUILD_WIDE_PAIRS=$wp timeout 300 python tools/build_bench.py --flags $fl --sizes 2000000 --repeat 8 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['case'], d['triangles'], 'ms', round(d['build_ms_median'],3), 'nodes', d['nodes'], end='  ')
print()"
    done
  done
done | tee $O/y_narrow_max.txt
