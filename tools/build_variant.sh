#!/bin/bash
# A/B builds of librtf_b200.so: tools/build_variant.sh NAME [nvcc flags...]  ->  raytracerfacility_b200/variants/librtf_NAME.so
# (git-ignored, travels to the GPU box; select with RTF_B200_LIB=...).  SRC=<dir> builds another source tree (e.g. a
# `git archive` of an older commit) with the same flags.
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
src=${SRC:-$root/raytracerfacility_b200/csrc}
obj=$root/build/variants/$name
mkdir -p "$obj" "$root/raytracerfacility_b200/variants"
for f in build render api materials prebake multi; do
  [ -f "$src/$f.cu" ] || continue
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC \
    -Xcompiler -fvisibility=hidden "$@" -c "$src/$f.cu" -o "$obj/$f.o" &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$root/raytracerfacility_b200/variants/librtf_$name.so" "$obj"/*.o -cudart static -lnccl $LINK_EXTRA
echo "built raytracerfacility_b200/variants/librtf_$name.so"
