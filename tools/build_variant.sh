#!/bin/bash
# Builds an A/B variant of the CUDA library: tools/build_variant.sh NAME -DGUAC_TK_AGG=1 ...
# -> build_variants/libguac_NAME.so (select it with GUAC_B200_LIB=build_variants/libguac_NAME.so)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p build_variants
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared "$@" \
  -o build_variants/libguac_$name.so guacamole_b200/csrc/guac_b200.cu
