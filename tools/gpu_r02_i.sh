#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/i_gpu_tests.log 2>&1; echo "gpu tests exit $?"; tail -3 $O/i_gpu_tests.log
VKRT_PARITY_EXACT=0 VKRT_B200_LIB=$PWD/vkrt_b200/csrc/libvkrt_b200_fmad.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -q -s -k "config1 or config3_full or config4_full or other_pose" > $O/i_fmad_tests.log 2>&1; echo "fmad tests exit $?"; grep -E "parity|passed|failed" $O/i_fmad_tests.log | tail -8
bash tools/gpu_r02_ab.sh i main fp32 fmad
