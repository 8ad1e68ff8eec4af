#!/bin/bash
# One bounded GPU session: new texture-LOD parity tests first, then the bench in both texture modes, then the rest of the GPU tier.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > gpurun_out/lod_gpu.txt 2>&1
timeout 150 python -m pytest tests/test_gpu_texture_lod.py -x -q -s > gpurun_out/lod_tests.log 2>&1; echo "lod tests exit $?" | tee -a gpurun_out/lod_tests.log
tail -5 gpurun_out/lod_tests.log
timeout 120 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/lod_bench_zero.json 2> gpurun_out/lod_bench_zero.err; echo "bench zero exit $?"
timeout 100 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --texture-lod cones > gpurun_out/lod_bench_cones.json 2> gpurun_out/lod_bench_cones.err; echo "bench cones exit $?"
cat gpurun_out/lod_bench_zero.json gpurun_out/lod_bench_cones.json | cut -c1-400
timeout 200 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/lod_gpu_parity.log 2>&1; echo "gpu parity exit $?" | tee -a gpurun_out/lod_gpu_parity.log
tail -3 gpurun_out/lod_gpu_parity.log
