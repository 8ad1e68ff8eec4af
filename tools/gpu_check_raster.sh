#!/bin/bash
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_raster.py -x -q -s > gpurun_out/raster_tests.log 2>&1; echo "raster tests exit $?" | tee -a gpurun_out/raster_tests.log
tail -8 gpurun_out/raster_tests.log
timeout 90 python tools/raster_bench.py --width 3840 --height 2160 > gpurun_out/raster_bench_4k.json 2> gpurun_out/raster_bench_4k.err; echo "bench 4k exit $?"
timeout 60 python tools/raster_bench.py --width 1280 --height 720 > gpurun_out/raster_bench_720p.json 2> gpurun_out/raster_bench_720p.err; echo "bench 720p exit $?"
cat gpurun_out/raster_bench_4k.json gpurun_out/raster_bench_720p.json
timeout 120 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_raster|k_mip' -c 40 --csv --log-file gpurun_out/raster_launches.csv python tools/raster_bench.py --width 3840 --height 2160 --steps 2 --warmup 1 > gpurun_out/raster_ncu.log 2>&1; echo "ncu exit $?"
tail -3 gpurun_out/raster_bench_4k.err
