#!/bin/bash
# Builds libdepthb200.so in-tree for sm_100a (cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
OUT=${OUT:-../libdepthb200.so}
SRCS="api.cu resample_plan.cu preprocess.cu postprocess.cu gemm_tcgen05.cu attention.cu attention_tc.cu attention_ws.cu misc.cu conv_direct.cu head_conv.cu $EXTRA_SRCS"
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared \
  -I../../include -I. ${NVCC_EXTRA} -o $OUT $SRCS -lcudart_static -ldl -lrt -lpthread 2>&1
echo "built $OUT"
