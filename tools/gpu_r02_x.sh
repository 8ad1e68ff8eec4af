#!/bin/bash
# 2-GPU check of the final state: the collected multi-GPU test, then bench.py at N=1 and N=2 on the same box
O=gpurun_out; mkdir -p $O
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_gpu_configs.py -m gpu -q -k multi_gpu > $O/x_multigpu_test.log 2>&1; echo "multi-GPU test exit $?"; tail -3 $O/x_multigpu_test.log
timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > $O/x_bench_n1.json 2> $O/x_bench_n1.err; echo "N=1 exit $?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > $O/x_bench_n2.json 2> $O/x_bench_n2.err; echo "N=2 exit $?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --workload raster --steps 10 --warmup 3 --no-cpu-baseline > $O/x_raster_n2.json 2> $O/x_raster_n2.err; echo "raster N=2 exit $?"
python - <<PY
import json
for n in (1,2):
    try:
        d=json.loads([l for l in open("$O/x_bench_n%d.json" % n) if l.startswith("{")][-1]); r=d["roofline"]
        print("N=%d" % n, round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"],3), "fnv", d["frame_fnv"], "stage", [round(x,2) for x in r["stage_ms_each"]])
    except Exception as e:
        print("N=%d failed" % n, e); print(open("$O/x_bench_n%d.err" % n).read()[-1500:])
try:
    d=json.loads([l for l in open("$O/x_raster_n2.json") if l.startswith("{")][-1]); print("raster N=2", round(d["value"]), d["ms_per_step"], d.get("frame_fnv"))
except Exception as e:
    print("raster failed", e); print(open("$O/x_raster_n2.err").read()[-800:])
PY
