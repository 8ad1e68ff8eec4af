#!/bin/bash
# robustness of the reinsertion passes: the whole GPU tier with the flag forced on every BLAS build, then the tier as it is
O=gpurun_out; mkdir -p $O
VKRT_BUILD_FAST_TRACE=1 timeout 2400 python -m pytest tests -m gpu -q --durations=5 > $O/aa_gpu_tier_forced.log 2>&1; echo "forced tier rc=$?"; tail -12 $O/aa_gpu_tier_forced.log
timeout 2400 python -m pytest tests -m gpu -x -q > $O/aa_gpu_tier.log 2>&1; echo "tier rc=$?"; tail -3 $O/aa_gpu_tier.log
