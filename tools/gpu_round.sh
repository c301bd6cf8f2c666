#!/bin/bash
# One GPU-box session: parity tests, smoke, bench, ncu launch list, ncu full capture of k_trace.
# usage (from the repo root, under gpurun): bash tools/gpu_round.sh <tag>
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log
python -c 'import __graft_entry__ as g; g.smoke()' > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $out/${tag}_smoke.log
python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?"; cat $out/${tag}_bench.json
bash tools/gpu_prof.sh $tag
ls -la $out | tail -12
