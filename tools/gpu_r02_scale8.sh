#!/bin/bash
# final scaling record on one 8-GPU box: N = 1, 4, 8, config 3, frame hashes must agree
O=gpurun_out; mkdir -p $O
echo "GPUs: $(nvidia-smi -L | wc -l)"
for n in 1 4 8; do
  if [ $n -eq 1 ]; then
    timeout 400 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline > $O/sc_bench_n$n.json 2> $O/sc_bench_n$n.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 20 --warmup 3 --no-cpu-baseline > $O/sc_bench_n$n.json 2> $O/sc_bench_n$n.err
  fi
  echo "bench N=$n exit $?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("$O/sc_bench_n$n.json") if l.startswith("{")][-1]); r=d["roofline"]
    print("N=$n", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"],3), "fnv", d["frame_fnv"], "stage", [round(x,2) for x in r["stage_ms_each"]])
except Exception as e:
    print("N=$n failed", e); print(open("$O/sc_bench_n$n.err").read()[-1200:])
PY
done
