#!/bin/bash
mkdir -p gpurun_out
timeout 60 python -m pytest tests/test_gpu_raster.py tests/test_gpu_texture_lod.py -x -q -s > gpurun_out/final_raster_lod_tests.log 2>&1; echo "raster+lod tests exit $?" | tee -a gpurun_out/final_raster_lod_tests.log
grep -E "passed|failed|raster vs|raster parity" gpurun_out/final_raster_lod_tests.log
VKRT_RASTER_BLOCKS=3 timeout 40 python tools/raster_bench.py --width 3840 --height 2160 > gpurun_out/final_raster_4k_b3.json 2>/dev/null; echo "b3 exit $?"
VKRT_RASTER_BLOCKS=4 timeout 40 python tools/raster_bench.py --width 3840 --height 2160 > gpurun_out/final_raster_4k_b4.json 2>/dev/null; echo "b4 exit $?"
cut -c1-200 gpurun_out/final_raster_4k_b3.json gpurun_out/final_raster_4k_b4.json
timeout 80 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/final_gpu_parity.log 2>&1; echo "parity exit $?" | tee -a gpurun_out/final_gpu_parity.log
tail -2 gpurun_out/final_gpu_parity.log
VKRT_RASTER_BLOCKS=4 timeout 40 python -m pytest tests/test_gpu_raster.py -x -q > gpurun_out/final_raster_b4_tests.log 2>&1; echo "raster b4 tests exit $?" | tee -a gpurun_out/final_raster_b4_tests.log
timeout 30 python tools/raster_bench.py --width 1280 --height 720 > gpurun_out/final_raster_720p.json 2>/dev/null; cut -c1-200 gpurun_out/final_raster_720p.json
