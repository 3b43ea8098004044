#!/bin/bash
# A/B of the node forms (float 256 B vs 8-bit 80 B), register caps and the SM-local tile remap on the C2 (708) and 10 M (2237) height fields.
mkdir -p gpurun_out
L=$PWD/radiance-ray-tracing_b200
for n in 708 2237; do
  echo "== q8 nodes, n=$n"
  RADIANCE_B200_LIB=$L/libradiance_b200_q8.so RD_EXP_VARIANTS='[{"flags":""},{"flags":"-DRD_FAST_MT=1"},{"flags":"-DRD_SM_LOCAL=7"},{"flags":"-DRD_SM_LOCAL=4"},{"flags":"-DRD_SM_LOCAL=14"},{"flags":"--maxrregcount=64"}]' \
    timeout 600 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH"
  echo "== f32 nodes, n=$n"
  RD_EXP_VARIANTS='[{"flags":"--maxrregcount=72"},{"flags":"--maxrregcount=72 -DRD_SM_LOCAL=7"},{"flags":""}]' \
    timeout 600 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH"
done
