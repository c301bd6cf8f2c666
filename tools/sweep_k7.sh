#!/bin/bash
# Dev sweep of the K7 wide walk's bookkeeping thresholds (262144 points of the config-2 scene).
run() { echo -n "$* : "; env "$@" python tools/dev_k7_bench.py --points 262144 --check 0 2>&1 | tail -1; }
run UAVMVS_K7_REFILL=20 UAVMVS_K7_STEPS=12
run UAVMVS_K7_REFILL=20 UAVMVS_K7_STEPS=24
run UAVMVS_K7_REFILL=24 UAVMVS_K7_STEPS=16
run UAVMVS_K7_REFILL=16 UAVMVS_K7_STEPS=16
run UAVMVS_K7_REFILL=20 UAVMVS_K7_STEPS=16 UAVMVS_K7_TRI=20
run UAVMVS_K7_REFILL=20 UAVMVS_K7_STEPS=16 UAVMVS_K7_TRI=28
