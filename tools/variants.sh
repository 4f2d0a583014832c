#!/bin/bash
# Tuning aid: time the stand-alone sort and one c2 frame bench under each libsplat_cuda_<name>.so given.
# usage: tools/variants.sh name1 name2 ...   ("default" = libsplat_cuda.so)
for v in "$@"; do
  if [ "$v" = default ]; then unset SPLAT_CUDA_LIB; else export SPLAT_CUDA_LIB=$PWD/gaussian_splat_b200/libsplat_cuda_$v.so; fi
  echo "== $v"
  python tools/sort_bench.py 67108864 64 5 64
  python tools/sort_bench.py 67108864 32 5 32
  python tools/sort_bench.py 12400000 32 10 32
  python tools/sort_bench.py 1583904 32 20 32
  python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-sort-bench | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c2 fps', round(d['value'],1), {k: round(v['ms']*1e3,1) for k,v in d['stages'].items()})"
done
