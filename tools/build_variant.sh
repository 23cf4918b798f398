#!/bin/bash
# Builds an experimental variant of libceracuda.so with extra nvcc -D flags (kernel experiments; select it with CERACUDA_LIB).
# usage: tools/build_variant.sh NAME -DFOO=1 ...   ->  build/var_NAME/libceracuda.so
set -e
NAME=$1; shift
D=build/var_$NAME; mkdir -p $D
SRC=ceracoder_b200/csrc
for f in cc_api cc_bitstream cc_cavlc cc_rc; do gcc -O2 -fPIC -std=gnu11 -I/usr/local/cuda/include -c $SRC/$f.c -o $D/$f.o; done
for f in cc_k_frame cc_k_motion cc_k_mb cc_k_cavlc cc_k_prims; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-O2 -Xptxas -v "$@" -c $SRC/$f.cu -o $D/$f.o 2> $D/$f.ptxas.log &
done
wait
/usr/local/cuda/bin/nvcc -shared -o $D/libceracuda.so $D/*.o -lm -lpthread 2>/dev/null
grep -A2 "k_deblock\|k_mb_encodeILi8" $D/cc_k_mb.ptxas.log | grep -E "Used|spill" | sed 's/ptxas info    : //'
