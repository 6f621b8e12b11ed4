#!/bin/bash
# Builds libdepthb200.so in-tree for sm_100a (cross-compiles without a GPU).  One object per translation unit,
# compiled in parallel into build/ (git-ignored); only sources newer than their object are recompiled.
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
OUT=${OUT:-../libdepthb200.so}
OBJ=${OBJ:-build}
SRCS="api.cu resample_plan.cu preprocess.cu postprocess.cu gemm_tcgen05.cu attention_ws.cu attention_ws2.cu misc.cu conv_direct.cu depthwise.cu head_conv.cu mbconv.cu engine.cu engine_nets.cu $EXTRA_SRCS"
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -I../../include -I. ${NVCC_EXTRA}"
mkdir -p $OBJ
echo "$FLAGS" > $OBJ/.flags.new
if ! cmp -s $OBJ/.flags.new $OBJ/.flags 2>/dev/null; then rm -f $OBJ/*.o; mv $OBJ/.flags.new $OBJ/.flags; fi
pids=()
objs=""
for s in $SRCS; do
  o=$OBJ/$(basename ${s%.*}).o
  objs="$objs $o"
  stale=0
  if [ ! -f $o ] || [ $s -nt $o ]; then stale=1; fi
  for h in *.h *.cuh ../../include/*.h; do if [ $h -nt $o ]; then stale=1; fi; done
  if [ $stale = 1 ]; then
    ( $NVCC $FLAGS -c -o $o $s 2>&1 || { rm -f $o; exit 1; } ) &
    pids+=($!)
  fi
done
rc=0
for p in "${pids[@]}"; do wait $p || rc=1; done
if [ $rc != 0 ]; then echo "nvcc compile failed"; exit 1; fi
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -Xcompiler -fPIC -o $OUT $objs -lcudart_static -ldl -lrt -lpthread 2>&1
echo "built $OUT"
# the no-Python host of the handle-level ABI (tests/c/engine_demo.c): plain C, linked against the library
gcc -O2 -std=c99 -I../../include -I/usr/local/cuda/include -o $OBJ/engine_demo ../../tests/c/engine_demo.c \
  -L.. -l:libdepthb200.so -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,'$ORIGIN/../..' 2>&1
echo "built $OBJ/engine_demo"
