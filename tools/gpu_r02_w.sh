#!/bin/bash
# where the FAST_TRACE build spends its time: launch list of one warm rebuild of the Sponza stand-in
O=gpurun_out; mkdir -p $O
timeout 300 python tools/build_bench.py --flags 8 --sizes 2000000 --repeat 3 > $O/w_build_ft.jsonl 2>&1; tail -2 $O/w_build_ft.jsonl | cut -c1-220
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_reinsert' --launch-skip 433 --launch-count 433 --csv --log-file $O/w_reinsert_launches.csv \
    python tools/build_bench.py --flags 8 --sizes 1000 --repeat 1 > $O/w_ncu.log 2>&1; echo "ncu exit $?"
python tools/ncu_summary.py launches $O/w_reinsert_launches.csv | head -12
