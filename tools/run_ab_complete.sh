#!/bin/bash
# experiment: complete 8-ary tree over the Morton order (RD_BUILD_COMPLETE = primitives per leaf slot) against the SAH tree
export RD_EXP_VARIANTS='[{"flags":""}]'
for n in 2237 708; do
  echo "== n=$n default"; timeout 300 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH" | tail -2
  for per in 1 2 3 4; do
    echo "== n=$n RD_BUILD_COMPLETE=$per"
    RD_BUILD_COMPLETE=$per timeout 300 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH" | tail -2
  done
done
