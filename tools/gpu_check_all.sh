#!/bin/bash
# Full GPU tier, rasteriser timing with and without the front-to-back order, one ncu capture of the rasteriser's top kernels.
mkdir -p gpurun_out
timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/all_gpu_tests.log 2>&1; echo "gpu tests exit $?" | tee -a gpurun_out/all_gpu_tests.log
tail -4 gpurun_out/all_gpu_tests.log
timeout 60 python tools/raster_bench.py --width 3840 --height 2160 > gpurun_out/raster_bench_4k_sorted.json 2> gpurun_out/raster_bench_4k_sorted.err; echo "bench sorted exit $?"
VKRT_RASTER_SORT=0 timeout 60 python tools/raster_bench.py --width 3840 --height 2160 > gpurun_out/raster_bench_4k_unsorted.json 2> gpurun_out/raster_bench_4k_unsorted.err; echo "bench unsorted exit $?"
timeout 60 python tools/raster_bench.py --width 1280 --height 720 > gpurun_out/raster_bench_720p_sorted.json 2>/dev/null
cut -c1-330 gpurun_out/raster_bench_4k_sorted.json gpurun_out/raster_bench_4k_unsorted.json gpurun_out/raster_bench_720p_sorted.json
timeout 100 ncu --set full --import-source on --clock-control none -k regex:'k_raster_large|k_raster_resolve|k_raster_small' -s 3 -c 3 -o gpurun_out/raster_full python tools/raster_bench.py --width 3840 --height 2160 --steps 1 --warmup 1 > gpurun_out/raster_ncu_full.log 2>&1; echo "ncu full exit $?"
timeout 60 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_raster|k_mip|k_onesweep' -c 40 --csv --log-file gpurun_out/raster_launches_v2.csv python tools/raster_bench.py --width 3840 --height 2160 --steps 1 --warmup 1 > gpurun_out/raster_ncu2.log 2>&1; echo "ncu list exit $?"
