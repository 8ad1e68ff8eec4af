#!/bin/bash
# reinsertion passes (VKRT_BUILD_PREFER_FAST_TRACE) A/B on one box: config 3 with the default build and with several schedules,
# then the builder / parity tests with the flag forced on
O=gpurun_out; mkdir -p $O
run() { # tag, env...
  local tag=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu-baseline --steps 6 > $O/t_${tag}.json 2> $O/t_${tag}.err || tail -3 $O/t_${tag}.err
  python - <<PY
import json
d=json.load(open("$O/t_${tag}.json")); r=d["roofline"]; b=d.get("bvh_build",{})
print("$tag", round(d["value"]), "ms", round(d["ms_per_step"],2), "fnv", d["frame_fnv"], "stage", [round(x,2) for x in r["stage_ms_each"][:8]], "any n/t", round(r["n_node_any"],2), round(r["n_tri_any"],2), "build ms", round(b.get("ms",0),2), "sah", round(b.get("sah_cost",0),2))
PY
}
run base A=1
run ft_default VKRT_BUILD_FAST_TRACE=1
run ft_3sweeps VKRT_BUILD_FAST_TRACE=1 VKRT_BUILD_REINSERT=3,12,3,4
run ft_plain16 VKRT_BUILD_FAST_TRACE=1 VKRT_BUILD_REINSERT=1,1,16,4
run ft_b10i2 VKRT_BUILD_FAST_TRACE=1 VKRT_BUILD_REINSERT=2,10,2,4
run ft_b12i5 VKRT_BUILD_FAST_TRACE=1 VKRT_BUILD_REINSERT=2,12,5,4
run base2 A=1
VKRT_BUILD_FAST_TRACE=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tlas.py -m gpu -x -q 2>&1 | tail -3
