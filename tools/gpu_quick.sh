#!/bin/bash
# Quick GPU check of a K1 change: the parity tests that cover K1 + two short bench lines (with / without an env switch).
# usage: bash tools/gpu_quick.sh <tag> [ENV=1 ...]
tag=${1:-rXX}; shift
mkdir -p gpurun_out
python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py tests/test_refcuda_gpu.py -m gpu -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${tag}_pytest.log
B='--steps 20 --warmup 5 --no-cpu-baseline --no-strong --no-extras'
python bench.py $B > gpurun_out/${tag}_bench_a.json 2>/dev/null; echo "bench a exit $?"
[ $# -gt 0 ] && { env "$@" python bench.py $B > gpurun_out/${tag}_bench_b.json 2>/dev/null; echo "bench b ($@) exit $?"; }
python - $tag <<'PY'
import json, sys
for s in "ab":
    try:
        d = json.load(open("gpurun_out/%s_bench_%s.json" % (sys.argv[1], s)))
        k = d["roofline"]["k1_kernels_ms"]
        print(s, "step %.3f e2e %.3f K1 %.3f masks %.3f cull %.3f trace %.3f retire %.3f K3 %.3f parity %s launches %d" % (
            d["ms_per_step"], d["e2e"]["ms_per_step"], d["stage_ms"]["K1"], k["k_view_masks"], k["k_cull_emit_q"], k["k_trace"], k["k_retire"],
            d["stage_ms"]["K3_K8"], d["parity_check"]["ok"], d["gpu_launches"]))
    except Exception as e:
        print(s, "n/a", e)
PY
