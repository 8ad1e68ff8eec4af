#!/bin/bash
# Scaling check on one box: N = 1, 2, 4, 8 (as many as the box has), frame hashes must agree; the collected multi-GPU test.
O=gpurun_out; mkdir -p $O
TAG=${1:-s}
NG=$(nvidia-smi -L | wc -l)
echo "GPUs: $NG"
timeout 900 python -m pytest tests/test_gpu_configs.py -m gpu -q -k multi_gpu > $O/${TAG}_multigpu_test.log 2>&1; echo "multi-GPU test exit $?"; tail -3 $O/${TAG}_multigpu_test.log
for n in 1 2 4 8; do
  if [ $n -le $NG ]; then
    if [ $n -eq 1 ]; then
      timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_n$n.json 2> $O/${TAG}_bench_n$n.err
    else
      timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_n$n.json 2> $O/${TAG}_bench_n$n.err
    fi
    echo "bench N=$n exit $?"
    python - <<PY
import json
try:
    d=json.loads([l for l in open("$O/${TAG}_bench_n$n.json") if l.startswith("{")][-1]); r=d["roofline"]
    print("N=$n", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"],3), "fnv", d["frame_fnv"], "ingest", {k: round(v,2) for k,v in d["ingest"].items()}, "stage", [round(x,2) for x in r["stage_ms_each"]])
except Exception as e:
    print("N=$n failed", e); print(open("$O/${TAG}_bench_n$n.err").read()[-1500:])
PY
  fi
done
