#!/bin/bash
# developer A/B build: tools/ab_build.sh <name> [extra nvcc flags...] -> build/lib_<name>.so  (select with THORMANG_B200_LIB)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC -shared -cudart static "$@" \
  -o build/lib_$name.so thormang_isaacgym_b200/csrc/tk_*.cu 2>&1 | grep -v "^$" | grep -iE "error|spill|registers" || true
echo build/lib_$name.so
