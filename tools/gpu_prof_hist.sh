#!/bin/bash
# ncu evidence for the spherical-histogram path (K4/K5/K6 batched and per position) and K7:
# launch list + one full capture per kernel.  usage (under gpurun): bash tools/gpu_prof_hist.sh <tag>
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
K='regex:^k_(view_masks|block_caps|sum_caps|cull_emit_q|trace|retire_hist|populate_spherical_histogram|evaluate_spherical_histogram|evaluate_spherical_histogram_batch|frustum_bits|estimate_capture_difficulty)'
python tools/dev_hist_bench.py --batch 1024 --single 8 > $out/${tag}_hist_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file $out/${tag}_hist_launches.csv \
    python tools/dev_hist_bench.py --batch 1024 --single 8 > $out/${tag}_hist_ncu_launches.log 2>&1
echo "hist launches exit $?"
# third batched pass (warm) onwards: skip the first two passes' launches (7 launches per pass, plus frustum bits once)
ncu --set full --clock-control none --import-source on -k "$K" -s 15 -c 12 -f -o $out/${tag}_hist \
    python tools/dev_hist_bench.py --batch 1024 --single 2 > $out/${tag}_hist_ncu_full.log 2>&1
echo "hist full exit $?"
python tools/dev_k7_bench.py --points 100000 --check 0 > $out/${tag}_k7_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:^k_estimate_capture_difficulty -s 1 -c 1 -f -o $out/${tag}_k7 \
    python tools/dev_k7_bench.py --points 100000 --check 0 > $out/${tag}_k7_ncu_full.log 2>&1
echo "k7 full exit $?"
tail -2 $out/${tag}_hist_plain.log $out/${tag}_k7_plain.log
