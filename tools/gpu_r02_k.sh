#!/bin/bash
O=gpurun_out; mkdir -p $O
TAG=${1:-k}
timeout 900 python -m pytest tests -m gpu -x -q -k "builder or rebuild or config5 or soup or collapse or device_input or radix" > $O/${TAG}_build_tests.log 2>&1; echo "builder tests exit $?"; tail -3 $O/${TAG}_build_tests.log
timeout 300 python tools/build_bench.py --flags 1 --sizes 2000000,10000000,50000000 2>&1 | tail -4 | tee $O/${TAG}_build_lbvh.jsonl
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file $O/${TAG}_build50M_lbvh.csv \
    python tools/build_bench.py --flags 1 --sizes 50000000 --repeat 1 --no-sponza > $O/${TAG}_build50M_lbvh.log 2>&1; echo "ncu build exit $?"
python tools/ncu_launch_bytes.py $O/${TAG}_build50M_lbvh.csv 2 | head -12
