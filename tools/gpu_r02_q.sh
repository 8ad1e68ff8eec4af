#!/bin/bash
O=gpurun_out; mkdir -p $O
echo skip
for v in main onefence acqrel; do
  if [ $v = main ]; then L=$PWD/vkrt_b200/csrc/libvkrt_b200.so; else L=$PWD/vkrt_b200/csrc/libvkrt_b200_$v.so; fi
  VKRT_B200_LIB=$L timeout 300 python tools/build_bench.py --flags 1 --sizes 10000000,50000000 2>&1 | tail -3 | cut -c1-200
  VKRT_B200_LIB=$L timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "builder or rebuild or device_input" 2>&1 | tail -1
done
