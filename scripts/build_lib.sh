#!/bin/bash
# Builds the C-ABI library in-tree (same flags as __graft_entry__.build) and prints the resource use of the assign kernels.
cd "$(dirname "$0")/.."
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared -Xptxas -v -Iinclude \
  -o flink-ml_b200/lib/libflinkml_kmeans_cuda.so flink-ml_b200/csrc/fmlk_api.cu -ldl 2>&1 | grep -A2 "assign_kernel\|error" | grep "spill\|Used\|error"
cuobjdump -sass -fun assign_kernel flink-ml_b200/lib/libflinkml_kmeans_cuda.so 2>/dev/null | awk '/Function :/{name=$3} /^\s+\/\*[0-9a-f]{4}\*\//{n[name]++} END{for(k in n) print n[k], "instrs", k}'
