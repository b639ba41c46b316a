#!/bin/bash
# Build libb200sim.so (C ABI + sm_100a kernels) in-tree.  Usage: ./build.sh [extra nvcc flags]
set -e
cd "$(dirname "$0")/asimov-mjlab_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $*"
mkdir -p ../_build
pids=()
for f in b2s_api k_smooth k_collision k_constraint k_solver k_post k_mdp k_mdp_tracking; do
  $NVCC $FLAGS -c $f.cu -o ../_build/$f.o &
  pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o ../libb200sim.so ../_build/*.o
echo "built $(cd .. && pwd)/libb200sim.so"
