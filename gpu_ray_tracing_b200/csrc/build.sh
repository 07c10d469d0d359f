#!/bin/sh
# Builds librt_b200.so (sm_100a only) next to the package.  -fmad=false: see rt_kernels.cu.
set -e
here="$(cd "$(dirname "$0")" && pwd)"
out="${RT_OUT:-$here/../librt_b200.so}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
HOSTCXX=/usr/bin/g++
[ -x "$HOSTCXX" ] || HOSTCXX=g++
"$NVCC" -ccbin "$HOSTCXX" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false $RT_EXTRA_NVCC_FLAGS \
    -Xcompiler -fPIC,-fvisibility=hidden,-ffp-contract=off,-Wall \
    -shared -o "$out" \
    "$here/rt_kernels.cu" "$here/rt_api.cu" "$here/rt_multi.cu" "$here/rt_host.cpp" \
    -lcudart_static -lpthread -ldl -lrt
echo "built $out"
