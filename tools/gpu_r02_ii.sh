#!/bin/bash
# graph-replayed fast rebuilds: tests, then build times with and without the graph
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tlas.py tests/test_gpu_raster.py -m gpu -x -q 2>&1 | tail -3
for g in graph nograph; do
  if [ $g = nograph ]; then E="VKRT_BUILD_NO_GRAPH=1"; else E="A=1"; fi
  echo "== $g"; env $E timeout 300 python tools/build_bench.py --flags 1 --sizes 262144,2000000,10000000 --repeat 10 2>&1 | cut -c1-175
done
