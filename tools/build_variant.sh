#!/bin/bash
# tools/build_variant.sh <name> <extra nvcc flags...>: a copy of libb200jpeg.so compiled with extra -D switches into
# tools/_variants/<name>/libb200jpeg.so (picked up with B200JPEG_LIB=...); kernel-variant experiments only.
name=$1; shift
dst=tools/_variants/$name
mkdir -p $dst/obj
C=nvimagecodec_b200/csrc
for f in decode_kernels huffman_kernels color_kernel encode_kernels; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden "$@" -c $C/$f.cu -o $dst/obj/$f.o &
done
wait
g++ -shared -o $dst/libb200jpeg.so $dst/obj/*.o $C/_obj/enc_plan.o $C/_obj/decoder.o $C/_obj/encoder.o $C/_obj/capi.o $C/_obj/jpeg_parse.o $C/_obj/batch_plan.o $C/_obj/log.o \
  -L/usr/local/cuda/lib64 -lcudart_static -ldl -lpthread -lrt && echo built $dst/libb200jpeg.so
