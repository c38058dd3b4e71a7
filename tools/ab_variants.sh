#!/bin/bash
# A/B timing of kernel variants on the GPU box: one short device-resident bench.py run (C2, 10 steps) per library
# build and per setting of THETAL_B200_FUSE_DSS, one line each with ms per step and the per-family kernel times.
#   variant A = the in-tree build (e3sm_b200/csrc/libthetal_b200.so);
#   variants B, C, ... = alternative builds placed in e3sm_b200/csrc/variants/lib<X>.so (git-ignored; they travel
#   with gpurun), e.g.  nvcc <Makefile flags> -DDSS_MINB=3 ... -o variants/libB.so
# usage (from the repository root):  VARIANTS="A B" FUSE="0 1" bash tools/ab_variants.sh
mkdir -p gpurun_out
for v in ${VARIANTS:-A}; do
  if [ $v = A ]; then unset THETAL_B200_LIB; else export THETAL_B200_LIB=$PWD/e3sm_b200/csrc/variants/lib$v.so; fi
  for f in ${FUSE:-1}; do
    THETAL_B200_FUSE_DSS=$f python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-strong > gpurun_out/ab_$v$f.json 2> gpurun_out/ab_$v$f.err
    python -c "
import json,sys
j=json.loads(open('gpurun_out/ab_$v$f.json').read().strip().splitlines()[-1]); k=j['kernel_ms_per_step']
print('$v fuse=$f', round(j['ms_per_step'],4), {a:round(b,3) for a,b in k.items() if b>0})"
  done
done
