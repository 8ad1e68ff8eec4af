#!/bin/bash
# acq_rel arrival counters in the cost-optimal collapse and the reinsertion refit: builder tests, build times, then the counters
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tlas.py -m gpu -x -q 2>&1 | tail -2
for fl in 0 8; do timeout 300 python tools/build_bench.py --flags $fl --sizes 2000000 --repeat 8 2>&1 | cut -c1-150; done
bash tools/gpu_r02_traffic.sh
