#!/bin/bash
# One GPU-box session: parity tests, smoke, bench, ncu launch list, ncu full capture of k_trace.
# usage (from the repo root, under gpurun): bash tools/gpu_round.sh <tag>
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log
python -c 'import __graft_entry__ as g; g.smoke()' > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $out/${tag}_smoke.log
python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?"; cat $out/${tag}_bench.json
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $out/${tag}_b2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $out/${tag}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on -k "regex:^k_trace$" -s 2 -c 1 -f -o $out/${tag}_k_trace \
    python tools/dev_bench.py --ploc --iters 4 > $out/${tag}_ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $out | tail -12
