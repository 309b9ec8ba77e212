#!/bin/bash
# tools/build_variant.sh <name> <extra -D flags...>: a copy of libb200jpeg.so compiled with extra -D switches into
# tools/_variants/<name>/libb200jpeg.so (picked up with B200JPEG_LIB=...); kernel-variant experiments only.
name=$1; shift
dst=tools/_variants/$name
mkdir -p $dst/obj
C=nvimagecodec_b200/csrc
for f in decode_kernels huffman_kernels color_kernel encode_kernels progressive_kernel noalloc_unstuff; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden "$@" -c $C/$f.cu -o $dst/obj/$f.o &
done
for f in enc_plan decoder encoder capi jpeg_parse batch_plan log; do
  g++ -O2 -std=c++17 -fPIC -fvisibility=hidden -Wno-unknown-pragmas -I/usr/local/cuda/include "$@" -c $C/$f.cpp -o $dst/obj/$f.o &
done
wait
g++ -shared -o $dst/libb200jpeg.so $dst/obj/*.o -L/usr/local/cuda/lib64 -lcudart_static -ldl -lpthread -lrt && echo built $dst/libb200jpeg.so
