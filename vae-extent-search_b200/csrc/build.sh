#!/bin/bash
# Builds libvaex_b200.so in-tree for sm_100a (cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
OUT=../libvaex_b200.so
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC"
mkdir -p build
pids=()
for f in abi simt_kernels elementwise topk segsum_tc head_tc chain_tc chain_tc3 chain_mc3 reshead3 train feature_blob kcenter sketch_features extent_text ascent_cluster ascent_compose collective; do
  if [ ! -f build/$f.o ] || [ $f.cu -nt build/$f.o ] || [ -n "$(find . -maxdepth 1 \( -name '*.cuh' -o -name '*.inc' \) -newer build/$f.o)" ] || [ ../../include/vaex_b200.h -nt build/$f.o ]; then
    $NVCC $FLAGS ${VAEX_PTXAS_V:+-Xptxas -v} -c $f.cu -o build/$f.o &
    pids+=($!)
  fi
done
if [ ! -f build/host_pack.o ] || [ host_pack.cpp -nt build/host_pack.o ]; then
  ${CXX:-g++} -O3 -std=c++17 -fPIC -mavx2 -mf16c -c host_pack.cpp -o build/host_pack.o &
  pids+=($!)
fi
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a build/*.o -o $OUT -lcuda -ldl
echo "built $(readlink -f $OUT)"
