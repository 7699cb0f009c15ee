#!/bin/sh
# Builds a variant of libssb200.so with extra -D flags for ONE source file (tuning aid):
#   tools/build_variant.sh <name> <source.cu> "<flags>"   -> tools/_bin/libssb200_<name>.so
# Run semantic-segmentation-server_b200/build.py first (the other objects are taken from its _obj/).
set -e
root="$(cd "$(dirname "$0")/.." && pwd)"
pkg="$root/semantic-segmentation-server_b200"
name="$1"; src="$2"; flags="$3"
mkdir -p "$root/tools/_bin"
obj="$root/tools/_bin/${src%.cu}_$name.o"
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I"$root/include" -I"$pkg/csrc" \
  -Xcompiler -fPIC,-fvisibility=hidden -diag-suppress 177 $flags -c "$pkg/csrc/$src" -o "$obj"
others=$(ls "$pkg"/csrc/_obj/*.o | grep -v "/${src%.cu}.o")
/usr/local/cuda/bin/nvcc -shared -o "$root/tools/_bin/libssb200_$name.so" $obj $others -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC
echo "$root/tools/_bin/libssb200_$name.so"
