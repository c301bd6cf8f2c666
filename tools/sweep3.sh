#!/bin/bash
# usage: bash tools/sweep3.sh <tag> "ENV=.. ENV=.." "ENV=.." ...   (one short bench line per argument)
tag=$1; shift
out=gpurun_out/${tag}_sweep.txt; mkdir -p gpurun_out; : > $out
for envs in "$@"; do
  line=$(env $envs python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-strong --no-extras 2>/dev/null)
  python - "$envs" "$line" >> $out <<'PY'
import json, sys
label, line = sys.argv[1], sys.argv[2]
try:
    d = json.loads(line); k = d["roofline"]["k1_kernels_ms"]
    print("%-44s step %.3f e2e %.3f K1 %.3f masks %.3f cull %.3f trace %.3f retire %.3f K3 %.3f parity %s" % (label, d["ms_per_step"],
        d["e2e"]["ms_per_step"], d["stage_ms"]["K1"], k["k_view_masks"], k["k_cull_emit_q"], k["k_trace"], k["k_retire"], d["stage_ms"]["K3_K8"], d["parity_check"]["ok"]))
except Exception as e:
    print("%-44s FAILED %r %s" % (label, e, line[:200]))
PY
  tail -1 $out
done
