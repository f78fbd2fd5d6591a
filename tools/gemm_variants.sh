#!/bin/bash
# dev tool: build A/B variants of libplsb.so that differ only in the DMMA GEMM's tile depth / pipeline stages
set -e
cd "$(dirname "$0")/.."
make -j8 pls_b200/libplsb.so > /dev/null
mkdir -p build/var
for v in "32 3 256" "32 3 512" "16 4 512" "32 2 512" "16 3 256"; do
  set -- $v
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Iinclude -DPLSB_GM_BK=$1 -DPLSB_GM_STAGES=$2 -DPLSB_GM_NT=$3 -Xptxas -v \
       -c pls_b200/csrc/gemm.cu -o build/var/gemm_$1_$2_$3.o 2>&1 | grep -A1 "dmma_atb" | grep -o "Used [0-9]* registers" | head -2
  objs=$(ls build/*.o | grep -v "build/gemm.o")
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/var/libplsb_$1_$2_$3.so $objs build/var/gemm_$1_$2_$3.o -ldl
done
ls -la build/var/*.so
