#!/bin/bash
# ncu launch list of one evaluation step (tree build filtered out) + full captures of the step's kernels.
# usage: bash tools/gpu_prof.sh <tag>
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
K='regex:^k_(view_masks|block_caps|sum_caps|add_u64|shaft|cull_emit|trace|retire|sort_eval|reduce|reduce_init|pack_score|fold_score|calculate_func_recons|process_observation_rays|evaluate_observation_rays)'
B='--steps 2 --warmup 1 --no-cpu-baseline --no-strong --no-extras'
python bench.py $B > $out/${tag}_b2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file $out/${tag}_launches.csv \
    python bench.py $B > $out/${tag}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on -k "$K" -s 24 -c 12 -f -o $out/${tag}_step \
    python tools/dev_bench.py --gsah --iters 4 > $out/${tag}_ncu_full.log 2>&1
echo "ncu full exit $?"
tail -3 $out/${tag}_ncu_full.log
