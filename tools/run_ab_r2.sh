#!/bin/bash
# A/B of the warp-synchronous triangle-phase wait (RD_TRI_WAIT) on the C2 (708) and 10 M (2237) height fields.
mkdir -p gpurun_out
for n in 708 2237; do
  echo "== n=$n"
  RD_EXP_VARIANTS='[{"flags":""},{"flags":"-DRD_TRI_WAIT=1"},{"flags":"-DRD_TRI_WAIT=2"},{"flags":"-DRD_TRI_WAIT=4"},{"flags":"-DRD_TRI_WAIT=6"},{"flags":"-DRD_TRI_WAIT=8"},{"flags":"-DRD_TRI_WAIT=4 --maxrregcount=72"},{"flags":"-DRD_TRI_WAIT=4 -DRD_TRI_WAIT_MAX=2"},{"flags":"-DRD_TRI_WAIT=4 -DRD_TRI_WAIT_MAX=8"},{"flags":"-DRD_TRI_WAIT=8 -DRD_TRI_WAIT_MAX=16"},{"flags":"-DRD_TRI_WAIT=8 -DRD_TRI_WAIT_MAX=16 --maxrregcount=72"},{"flags":"-DRD_TRI_WAIT=4 -DRD_FAST_MT=1"}]' \
    timeout 600 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH"
done
