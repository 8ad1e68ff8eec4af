#!/bin/bash
# depth-1 launches: lane refill for the incoherent reflection rays, on one GPU at size 1 and as one rank of 8
O=gpurun_out; mkdir -p $O
for size in 1 8; do
  for rf in none 8 16 24; do
    if [ $rf = none ]; then E="A=1"; else E="VKRT_TRACE_REFILL_BOUNCE=$rf"; fi
    echo -n "refill $rf: "; env $E timeout 300 python tools/rank_stage_times.py --size $size 2>&1 | tail -1
  done
done | tee $O/v_refill_bounce.txt
