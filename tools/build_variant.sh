#!/bin/sh
# Developer tool: build a variant of libwenoext_b200.so with extra nvcc flags into build/variants/<name>/
# (used through WENOEXT_LIB=<path> for same-box A/B runs).   tools/build_variant.sh <name> [flags...]
set -e
cd "$(dirname "$0")/.."
name=$1; shift
out=build/variants/$name
mkdir -p $out
for f in kernels plan pinv noalloc ls_rows; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I include "$@" $( [ $f = ls_rows ] && echo --fmad=false ) \
       -c wenoext_b200/csrc/$f.cu -o $out/$f.o &
done
wait
nvcc -shared -o $out/libwenoext_b200.so $out/kernels.o $out/plan.o $out/pinv.o $out/noalloc.o $out/ls_rows.o -lpthread -ldl -cudart static
echo $out/libwenoext_b200.so
