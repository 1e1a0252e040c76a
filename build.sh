#!/usr/bin/env bash
# Builds the C-ABI library in-tree for sm_100a (nvcc cross-compiles without a GPU) and the oracle.
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
mkdir -p build   # ptxas log and scratch; git-ignored, so a fresh checkout has none
OUT=terraform-provider-iterative_b200/libtpimd5.so
$NVCC -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a \
      -Xcompiler -fPIC,-Wall,-Wextra,-fvisibility=hidden -Xptxas -v \
      -shared -o "$OUT" csrc/tpimd5.cu csrc/manifest.cc -lpthread 2> build/ptxas.log || { cat build/ptxas.log; exit 1; }
grep -E "error|warning" build/ptxas.log | grep -v "ptxas info" || true
g++ -std=c++17 -O2 -Wall -Wextra -o terraform-provider-iterative_b200/tpimd5sum csrc/tpimd5sum.cc \
    -Lterraform-provider-iterative_b200 -ltpimd5 -Wl,-rpath,'$ORIGIN'
make -s -C oracle
echo "built $OUT"
