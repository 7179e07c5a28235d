#!/bin/bash
# experiment builds of libccgpu with extra -D flags into build/variants/<name>.so:  tools/build_variant.sh name -DCCG_PF_UNROLL=4 ...
set -e
NAME=$1; shift
SRC=cloudclean_b200/csrc; OUT=build/variants/$NAME; mkdir -p $OUT
FL="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr $*"
for f in query; do /usr/local/cuda/bin/nvcc $FL -c $SRC/$f.cu -o $OUT/$f.o & done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/variants/$NAME.so $OUT/query.o $SRC/build.o $SRC/radsel.o $SRC/gridnormals.o $SRC/noalloc.o $SRC/voxel.o $SRC/ptx.o $SRC/api.o -lcudart_static -lpthread -ldl -lrt
ls -la build/variants/$NAME.so
