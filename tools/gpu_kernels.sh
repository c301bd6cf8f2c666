#!/bin/bash
# Per-kernel times of one evaluation step (ncu launch list over tools/dev_bench.py); dev tool for A/B runs.
# usage (under gpurun): bash tools/gpu_kernels.sh [dev_bench args]
K='regex:^k_(view_masks|block_caps|sum_caps|add_u64|shaft|entry|cull_emit|cull_emit_q|trace|retire|retire_hist|sort_eval|reduce|reduce_init)'
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 80 --csv --log-file gpurun_out/kl.csv \
    python tools/dev_bench.py --gsah --iters 4 "$@" > gpurun_out/kl.log 2>&1
python tools/ncu_launches.py gpurun_out/kl.csv | sed 's/per-step.*share/share/'
