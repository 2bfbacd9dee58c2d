#!/bin/bash
# on the GPU box: tools/run_variants.sh "<bench args>" v1 v2 ...  -> gpurun_out/var_<v>.json (one bench line each)
args=$1; shift
cp neural-renderer_b200/libnrb200.so /tmp/libnrb200_main.so
for v in "$@"; do
  cp variants/$v.so neural-renderer_b200/libnrb200.so
  python bench.py --no-cpu-baseline --no-gpu-baseline $args > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err
done
cp /tmp/libnrb200_main.so neural-renderer_b200/libnrb200.so
