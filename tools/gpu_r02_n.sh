#!/bin/bash
O=gpurun_out; mkdir -p $O
TAG=${1:-n}
timeout 600 python -m pytest tests/test_gpu_raster.py tests/test_gpu_zz_host_rasterizer.py -m gpu -x -q > $O/${TAG}_raster_tests.log 2>&1; echo "raster tests exit $?"; tail -2 $O/${TAG}_raster_tests.log
for rep in 1 2; do
timeout 300 python bench.py --workload raster --steps 20 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_raster.json 2> $O/${TAG}_bench_raster.err
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench_raster.json"))
print("raster", round(d["value"]), d["unit"], "ms", round(d["ms_per_step"],3), "fnv", d["frame_fnv"])
PY
VKRT_RASTER_SORT=1 timeout 300 python bench.py --workload raster --steps 20 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_raster_sorted.json 2> $O/${TAG}_bench_raster_sorted.err
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench_raster_sorted.json"))
print("raster sorted", round(d["value"]), d["unit"], "ms", round(d["ms_per_step"],3), "fnv", d["frame_fnv"])
PY
done
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_raster' --launch-skip 28 --launch-count 14 --csv --log-file $O/${TAG}_raster_launches.csv python bench.py --workload raster --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_raster_ncu.log 2>&1; echo "ncu raster exit $?"
