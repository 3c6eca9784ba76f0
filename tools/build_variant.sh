#!/bin/bash
# tools/build_variant.sh NAME "-DFLAG ..." : builds lib/var_NAME.so with extra nvcc flags (kernel experiments)
set -e
cd "$(dirname "$0")/../3d-gaussian-splat-attack_b200/csrc"
NAME=$1; shift
B=../build_var_$NAME
mkdir -p $B ../lib
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC --expt-extended-lambda $*"
for f in abi static_scratch preprocess_fwd binning radix_sort blend_fwd blend_bwd preprocess_bwd preprocess_bwd_views attack_ops; do
  nvcc $FLAGS -c -o $B/$f.o $f.cu 2>&1 | grep -E "error|Used|spill|Compiling" &
done
wait
nvcc -shared -o ../lib/var_$NAME.so $B/*.o -gencode arch=compute_100a,code=sm_100a
rm -rf $B
echo built lib/var_$NAME.so
