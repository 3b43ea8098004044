#!/bin/bash
# user-program compile flags: square root at OpenCL accuracy (3 ulp; --prec-sqrt=false) and denormal flush (--ftz=true, allowed by
# OpenCL for single precision) on the primary frame (10 M, 1 M) and on the C3 path tracer (8 spp, inline + wavefront)
for n in 2237 708; do
  echo "== primary n=$n"
  RD_EXP_VARIANTS='[{"flags":""},{"flags":"--prec-sqrt=false"},{"flags":"--ftz=true"},{"flags":"--prec-sqrt=false --ftz=true"}]' timeout 300 python tools/exp_variants.py $n 2>&1 | grep -v "Max BVH\|^bvh"
done
for f in "" "--prec-sqrt=false" "--prec-sqrt=false --ftz=true"; do
  echo "== C3 8 spp flags='$f'"
  RD_NVRTC_FLAGS="$f" timeout 600 python tools/exp_path.py c3 2237 8 5 2>&1 | grep -v "Max BVH" | tail -4
done
