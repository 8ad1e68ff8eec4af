#!/bin/bash
# no-alloc entry point: alone in a fresh process, under memcheck, then the whole GPU tier (it is collected last), smoke(),
# and the builder lines with the acq_rel climb as default
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,memory.used --format=csv,noheader
timeout 900 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/s_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -3 $O/s_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q --durations=8 > $O/s_gpu_tier.log 2>&1; echo "tier rc=$?"; tail -14 $O/s_gpu_tier.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/s_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/s_smoke.log
timeout 600 python tools/build_bench.py --flags 1 --sizes 10000000,50000000 > $O/s_build.log 2>&1; tail -3 $O/s_build.log | cut -c1-300
