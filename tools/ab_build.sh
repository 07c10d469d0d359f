#!/bin/sh
# Developer A/B: builds librt_b200 variants with different -D settings into tools/ab/<name>.so
# usage: tools/ab_build.sh name "-DRT_PRIM_THREADS=352 -DRT_PRIM_CTAS=2"
set -e
here="$(cd "$(dirname "$0")" && pwd)"
mkdir -p "$here/ab"
RT_OUT="$here/ab/$1.so" RT_EXTRA_NVCC_FLAGS="$2" sh "$here/../gpu_ray_tracing_b200/csrc/build.sh"
