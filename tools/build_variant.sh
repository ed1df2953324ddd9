#!/bin/bash
# A/B builds while tuning: tools/build_variant.sh NAME [extra nvcc flags] -> build_exp/NAME/libtdod.so (use with TDOD_LIB=...)
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
out=$root/build_exp/$name
mkdir -p $out
for f in $root/3dod_b200/csrc/*.cu; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Wno-deprecated-gpu-targets "$@" -c $f -o $out/$(basename ${f%.cu}).o &
done
wait
/usr/local/cuda/bin/nvcc -shared -o $out/libtdod.so $out/*.o -lcudart -lpthread
rm -f $out/*.o
echo $out/libtdod.so
