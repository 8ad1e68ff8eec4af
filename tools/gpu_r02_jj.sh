#!/bin/bash
for nm in 256 128 64; do for wp in 12 8 6; do echo -n "narrowMax $nm pairs $wp: "; VKRT_BUILD_NARROW_MAX=$nm VKRT_BUILD_WIDE_PAIRS=$wp timeout 300 python tools/build_bench.py --flags 1 --sizes 2000000 --repeat 10 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['case'], 'ms', round(d['build_ms_median'],3), 'nodes', d['nodes'], end='  ')
print()"; done; done
