#!/bin/bash
# full GPU tier with the reinsertion tests, then every static workload on the new default build flavour
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -x -q --durations=5 > $O/u_gpu_tier.log 2>&1; echo "tier rc=$?"; tail -9 $O/u_gpu_tier.log
for w in config3 config1 config2 config4; do
  timeout 400 python bench.py --workload $w --no-cpu-baseline --steps 10 > $O/u_bench_$w.json 2> $O/u_bench_$w.err || tail -3 $O/u_bench_$w.err
  python - <<PY
import json
d=json.load(open("$O/u_bench_$w.json")); r=d["roofline"]; b=d["bvh_build"]
print("$w", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"],2), "fnv", d["frame_fnv"], "any n/t", round(r["n_node_any"],2), round(r["n_tri_any"],2), "closest n/t", round(r["n_node_closest"],2), round(r["n_tri_closest"],2), "build", b["flavour"], round(b["warm_ms"],2), "sah", round(b["sah_cost"],2))
PY
done
