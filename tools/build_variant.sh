#!/bin/sh
# Build a tuning variant of the product library: tools/build_variant.sh NAME "-DENTB_THREADS=512 ..."
# -> build/lib_NAME.so ; run with IMAGE_CODECS_B200_LIB=build/lib_NAME.so python bench.py ...
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
SRC=$ROOT/image_codecs_b200/csrc
mkdir -p "$ROOT/build"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden $2 \
	-c "$SRC/jpeg_cuda.cu" -o "$ROOT/build/jpeg_cuda_$1.o"
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$ROOT/build/lib_$1.so" "$ROOT/build/jpeg_cuda_$1.o" \
	"$SRC/jpeg_parse.o" "$SRC/image_api.o" -cudart static -lpthread
echo "$ROOT/build/lib_$1.so"
