#!/bin/bash
# Build a variant of libpvvote.so with extra nvcc flags into tools/_build/<name>.so (development A/B and debug builds).
# usage: tools/build_variant.sh <name> [nvcc flags...]      then: PVV_LIB_PATH=tools/_build/<name>.so python ...
set -e
cd "$(dirname "$0")/.."
NAME=$1; shift
OUT=tools/_build/$NAME
mkdir -p $OUT
for f in pvv_compact.cu pvv_vote.cu pvv_finalize.cu pvv_primitives.cu pvv_pnp.cu pvv_capi.cu pvv_host.cu pvv_host_simd.cpp; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC "$@" \
      -c pvnet_b200/csrc/$f -o $OUT/${f%.*}.o &
done
wait
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -cudart static $OUT/*.o -o tools/_build/$NAME.so
echo tools/_build/$NAME.so
