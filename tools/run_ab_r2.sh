#!/bin/bash
# A/B of L2 persistence for the node array (option l2_persist) on the 10 M (2237) and C2 (708) height fields.
mkdir -p gpurun_out
export RD_EXP_VARIANTS='[{"flags":""}]'
for n in 2237 708; do
  for mb in 0 32 64 96; do
    echo "== n=$n l2_persist=${mb} MB"
    RD_EXP_L2_PERSIST=$((mb*1024*1024)) timeout 600 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH\|^bvh"
  done
done
