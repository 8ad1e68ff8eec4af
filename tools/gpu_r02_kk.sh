#!/bin/bash
# after the graph-replayed fast rebuilds: full GPU tier, builder lines, counters for the new hash
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -x -q > $O/fin_gpu_tests.log 2>&1; echo "tier rc=$?"; tail -3 $O/fin_gpu_tests.log
timeout 300 python tools/build_bench.py --flags 1 > $O/fin_build_lbvh.jsonl 2>&1; cut -c1-160 $O/fin_build_lbvh.jsonl
timeout 400 python bench.py --workload config5 --steps 5 --warmup 3 --no-cpu-baseline > $O/fin_bench_config5.json 2> $O/fin_bench_config5.err; echo "config5 exit $?"
bash tools/gpu_r02_traffic.sh
