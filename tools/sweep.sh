#!/bin/bash
# dev: run tools/dev_bench.py against every library variant in build_variants/
for f in build_variants/lib_*.so; do
  n=$(basename $f .so)
  r=$(UAVMVS_B200_LIB=$PWD/$f python tools/dev_bench.py --gsah --iters 5 2>&1 | grep stages | tail -2 | tr '\n' ' ')
  echo "$n $r"
done
