#!/bin/bash
O=gpurun_out; mkdir -p $O
TAG=${1:-l}
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1; echo "gpu tests exit $?"; tail -4 $O/${TAG}_gpu_tests.log
timeout 300 python bench.py --no-cpu-baseline --steps 10 > $O/${TAG}_bench_config3.json 2> $O/${TAG}_bench_config3.err
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench_config3.json")); r=d["roofline"]
print("config3", round(d["value"]), "ms", round(d["ms_per_step"],2), "fnv", d["frame_fnv"], "stage", [round(x,2) for x in r["stage_ms_each"]])
PY
timeout 300 python bench.py --workload raster --steps 20 --warmup 3 > $O/${TAG}_bench_raster.json 2> $O/${TAG}_bench_raster.err; echo "raster bench exit $?"
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench_raster.json"))
print("raster", round(d["value"]), d["unit"], "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), "fnv", d["frame_fnv"], "cpu", d.get("cpu_baseline",{}).get("value"))
PY
NG=$(nvidia-smi -L | wc -l)
if [ $NG -ge 2 ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --workload raster --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_raster_n2.json 2> $O/${TAG}_bench_raster_n2.err; echo "raster N=2 exit $?"
  tail -1 $O/${TAG}_bench_raster_n2.json | cut -c1-400
fi
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_raster' --launch-skip 24 --launch-count 12 --csv --log-file $O/${TAG}_raster_launches.csv python bench.py --workload raster --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_raster_ncu.log 2>&1; echo "ncu raster exit $?"
python tools/ncu_launch_bytes.py $O/${TAG}_raster_launches.csv 2 | head -10
