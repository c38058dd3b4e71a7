for v in A; do
  if [ $v = A ]; then export -n THETAL_B200_LIB; unset THETAL_B200_LIB; else export THETAL_B200_LIB=$PWD/e3sm_b200/csrc/variants/lib$v.so; fi
  for f in 1; do
    THETAL_B200_FUSE_DSS=$f python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-strong > gpurun_out/ab_$v$f.json 2> gpurun_out/ab_$v$f.err
    python -c "
import json,sys
j=json.loads(open('gpurun_out/ab_$v$f.json').read().strip().splitlines()[-1]); k=j['kernel_ms_per_step']
print('$v fuse=$f', round(j['ms_per_step'],4), {a:round(b,3) for a,b in k.items() if b>0})"
  done
done
