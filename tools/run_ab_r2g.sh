#!/bin/bash
# A/B of build-quality knobs on the primary frame (10 M and 1 M height fields): SAH primitive cost, primitives per leaf slot,
# LBVH topology. Library variants are built with `make all EXTRA=... OUT=../libradiance_b200_<name>.so OBJDIR=obj_<name>`.
mkdir -p gpurun_out
export RD_EXP_VARIANTS='[{"flags":""}]'
HERE=$(cd "$(dirname "$0")/.." && pwd)
for n in 2237 708; do
  for v in default sah3 sah8 leaf1 leaf2 leaf4; do
    lib=$HERE/radiance-ray-tracing_b200/libradiance_b200_$v.so
    [ $v = default ] && lib=$HERE/radiance-ray-tracing_b200/libradiance_b200.so
    [ -f $lib ] || continue
    echo "== n=$n lib=$v"
    RADIANCE_B200_LIB=$lib timeout 300 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH" | tail -2
  done
  echo "== n=$n lib=default RD_BUILD_LBVH=1"
  RD_BUILD_LBVH=1 timeout 300 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH" | tail -2
done
