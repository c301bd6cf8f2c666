#!/bin/bash
# Dev sweep on the GPU box: library variants (build_variants/*.so) x traversal thresholds, short bench runs.
# usage: bash tools/sweep2.sh <tag>   -> gpurun_out/<tag>_sweep.txt
tag=${1:-rXX}
out=gpurun_out/${tag}_sweep.txt
mkdir -p gpurun_out
: > $out
run() {   # label, lib, env...
  label=$1; lib=$2; shift 2
  line=$(env "$@" UAVMVS_B200_LIB=$lib python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-strong --no-extras 2>/dev/null)
  python - "$label" "$line" >> $out <<'PY'
import json, sys
label, line = sys.argv[1], sys.argv[2]
try:
    d = json.loads(line)
    k = d["roofline"]["k1_kernels_ms"]
    print("%-28s step %.3f ms  e2e %.3f  K1 %.3f  cull %.3f trace %.3f retire %.3f  K3 %.3f  parity %s" % (
        label, d["ms_per_step"], d["e2e"]["ms_per_step"], d["stage_ms"]["K1"], k["k_cull_emit_q"], k["k_trace"], k["k_retire"],
        d["stage_ms"]["K3_K8"], d["parity_check"]["ok"]))
except Exception as e:
    print("%-28s FAILED %r %s" % (label, e, line[:200]))
PY
  tail -1 $out
}
D=$PWD/uavmvs_b200/libuavmvs_b200.so
run "default r20" $D A=1
run "default r16" $D UAVMVS_K1_REFILL=16
run "default r12" $D UAVMVS_K1_REFILL=12
run "default r8" $D UAVMVS_K1_REFILL=8
run "default r4" $D UAVMVS_K1_REFILL=4
run "cull-setup r20" $D UAVMVS_CULL_SETUP=1
run "default r12 tri8" $D UAVMVS_K1_REFILL=12 UAVMVS_K1_TRI=8
run "default r12 steps4" $D UAVMVS_K1_REFILL=12 UAVMVS_K1_STEPS=4
run "default r12 steps8" $D UAVMVS_K1_REFILL=12 UAVMVS_K1_STEPS=8
for v in prmt8 prmt7 prmt6 base7; do
  L=$PWD/build_variants/libuav_$v.so
  [ -f $L ] || continue
  run "$v r20" $L A=1
  run "$v r12" $L UAVMVS_K1_REFILL=12
done
