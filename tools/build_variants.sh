#!/bin/bash
# tools/build_variants.sh "name:-DFOO=1 -DBAR=0" ... — builds variants/libvectordb_<name>.so (timing experiments;
# select one with VECTORDB_NATIVE_PATH). Not part of the product build.
set -e
mkdir -p "$(dirname "$0")/../variants"
cd "$(dirname "$0")/../vector_db_b200/csrc"
SRCS="vdb_api.cu pq_kernels.cu adc_fast.cu hnsw_kernels.cu flat_kernels.cu flat_gemm.cu snapshot.cu host_graph.cpp"
for v in "$@"; do
  name="${v%%:*}"; flags="${v#*:}"
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-O3,-mavx2,-mfma $flags -shared \
       -o ../../variants/libvectordb_$name.so $SRCS 2> ../../variants/build_$name.log &
done
wait
ls -la ../../variants/*.so
