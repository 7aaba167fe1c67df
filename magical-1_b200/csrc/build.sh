#!/bin/sh
# Build libmagical_b200.so in-tree for sm_100a (B200).  -fmad=false keeps double arithmetic
# identical to the Chipmunk operation order (no FMA contraction).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -std=c++17 -Xcompiler -fPIC,-fno-fast-math,-ffp-contract=off"
$NVCC $FLAGS -shared -o "${OUT:-../libmagical_b200.so}" mg_api.cu mg_physics.cu mg_raster.cu mg_frames.cu mg_tables.cpp mg_reset.cpp "$@"
