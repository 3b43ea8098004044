#!/bin/bash
# north_star design points with a number: (a) short traversal stack in shared memory (RD_SMEM_STACK entries per lane, the rest
# thread-local), (b) what tighter child boxes could buy: fetches per ray with exact float boxes (the limit of 16-bit boxes)
# against the 8-bit boxes.
HERE=$(cd "$(dirname "$0")/.." && pwd)
for n in 2237 708; do
  echo "== n=$n 8-bit nodes, stack: thread-local / 4 / 8 / 12 entries in shared memory"
  RD_EXP_VARIANTS='[{"flags":""},{"flags":"-DRD_SMEM_STACK=4"},{"flags":"-DRD_SMEM_STACK=8"},{"flags":"-DRD_SMEM_STACK=12"}]' timeout 300 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH\|^bvh"
  echo "== n=$n float-box nodes (256 B)"
  RADIANCE_B200_LIB=$HERE/radiance-ray-tracing_b200/libradiance_b200_f32.so RD_EXP_VARIANTS='[{"flags":""}]' timeout 300 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH\|^bvh"
done
