#!/bin/bash
# Tuning aid: one c2 frame bench (stage times) under each libsplat_cuda_<name>.so given ("default" = libsplat_cuda.so).
for v in "$@"; do
  if [ "$v" = default ]; then unset SPLAT_CUDA_LIB; else export SPLAT_CUDA_LIB=$PWD/gaussian_splat_b200/libsplat_cuda_$v.so; fi
  echo "== $v"
  python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-sort-bench ${WL:+--workload $WL} | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('fps', round(d['value'],1), {k: round(v['ms']*1e3,1) for k,v in d['stages'].items()})"
done
