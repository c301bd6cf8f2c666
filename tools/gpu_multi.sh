#!/bin/bash
# Multi-GPU session (gpurun --gpus N): the world-2 NCCL test through the C ABI, then bench.py under torchrun at N ranks
# (weak config-2 line with the collective inside the timed call + the config-5 strong block), then the reference arm.
# usage: bash tools/gpu_multi.sh <tag> <N> [bench flags]
tag=${1:-rXX}; n=${2:-2}; shift 2
mkdir -p gpurun_out
python -m pytest tests/test_sharded_gpu.py -m gpu -q > gpurun_out/${tag}_pytest_sharded.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${tag}_pytest_sharded.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 20 --warmup 5 "$@" \
    > gpurun_out/${tag}_bench_n$n.json 2> gpurun_out/${tag}_bench_n$n.err; echo "bench N=$n exit $?"
tail -c 600 gpurun_out/${tag}_bench_n$n.err
python - gpurun_out/${tag}_bench_n$n.json <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print("weak: %.3f ms/step, %.3f Grays/s, e2e %.3f ms, collective %.3f ms, parity %s" % (d["ms_per_step"], d["value"] / 1e9, d["e2e"]["ms_per_step"], d["stage_ms"]["score_collective"], d["parity_check"]["ok"]))
print("per rank", d["per_rank"]["rows"])
if "strong" in d:
    s = d["strong"]
    print("strong: %.3f ms/step, %.3f Grays/s, parity %s" % (s["ms_per_step"], s["value"] / 1e9, s.get("parity_check", {}).get("ok")))
    print("per rank", s["per_rank"]["rows"])
PY
