#!/bin/bash
# one full ncu capture of the two dominant reinsertion kernels (a mid-size band of the first sweep) on the Sponza stand-in
O=gpurun_out; mkdir -p $O
timeout 600 ncu --set full --import-source on --clock-control none -k regex:'k_reinsert_(search|refit)' --launch-skip 56 --launch-count 4 -o $O/gg_reinsert_full -f \
    python tools/build_bench.py --flags 8 --sizes 1000 --repeat 1 > $O/gg_ncu.log 2>&1; echo "ncu exit $?"
ls -la $O/gg_reinsert_full.ncu-rep
