#!/bin/bash
# Build a variant library for kernel experiments: tools/build_variant.sh <name> [nvcc -D flags...]  ->  asimov-mjlab_b200/_variants/lib_<name>.so
set -e
name=$1; shift
cd "$(dirname "$0")/../asimov-mjlab_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $*"
out=../_variants/$name; mkdir -p $out
pids=()
for f in b2s_api k_smooth k_collision k_constraint k_solver k_post k_mdp k_mdp_tracking; do
  $NVCC $FLAGS -c $f.cu -o $out/$f.o & pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o ../_variants/lib_$name.so $out/*.o
rm -rf $out
echo "built lib_$name.so"
