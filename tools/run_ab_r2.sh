#!/bin/bash
# A/B of the node forms (float 256 B vs 8-bit 80 B) and register caps on the C2 (708) and 10 M (2237) height fields.
mkdir -p gpurun_out
L=$PWD/radiance-ray-tracing_b200
for n in 708 2237; do
  echo "== f32 nodes, n=$n"
  RD_EXP_VARIANTS='[{"flags":""},{"flags":"--maxrregcount=80"},{"flags":"--maxrregcount=72"},{"flags":"--maxrregcount=64"},{"flags":"-DRD_FAST_MT=1"},{"flags":"","block":64},{"flags":"","block":256}]' \
    timeout 600 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH"
  echo "== q8 nodes, n=$n"
  RADIANCE_B200_LIB=$L/libradiance_b200_q8.so RD_EXP_VARIANTS='[{"flags":""},{"flags":"-DRD_FAST_MT=1"}]' \
    timeout 600 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH"
done
