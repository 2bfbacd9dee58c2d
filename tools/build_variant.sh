#!/bin/bash
# usage: tools/build_variant.sh NAME "<extra nvcc flags>"  -> variants/NAME.so (git-ignored; travels with gpurun)
set -e
here=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p "$here/variants"
make -C "$here/neural-renderer_b200/csrc" -j8 BUILD="$here/variants/build_$1" OUT="$here/variants/$1.so" EXTRA="$2" 2>&1 | grep -E "error|ptxas info.*k_bounce_begin|registers" | head -20
ls -la "$here/variants/$1.so"
