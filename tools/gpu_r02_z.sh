#!/bin/bash
# launch list of warm Sponza-sized builds, fast build (flags 1) and default (flags 0)
O=gpurun_out; mkdir -p $O
for fl in 1 0; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 200 --launch-count 400 --csv --log-file $O/z_build_launches_f$fl.csv \
      python tools/build_bench.py --flags $fl --sizes 1000 --repeat 12 > $O/z_ncu_f$fl.log 2>&1; echo "ncu flags $fl exit $?"
  python tools/ncu_summary.py launches $O/z_build_launches_f$fl.csv | head -24
done
timeout 300 python tools/build_bench.py --flags 1 --sizes 2000000,10000000 --repeat 8 2>&1 | cut -c1-160
