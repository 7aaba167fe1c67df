#!/bin/sh
# Build libmagical_b200.so in-tree for sm_100a (B200).  -fmad=false keeps double arithmetic
# identical to the Chipmunk operation order (no FMA contraction).
# The translation units are compiled side by side (objects under csrc/_obj, git-ignored) and linked once.
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -std=c++17 -Xcompiler -fPIC,-fno-fast-math,-ffp-contract=off"
OBJ=${OBJ_DIR:-_obj}
mkdir -p "$OBJ"
SRCS="mg_physics.cu mg_raster.cu mg_api.cu mg_frames.cu mg_tables.cpp mg_reset.cpp mg_hostasm.cpp"
pids=""
for f in $SRCS; do
    $NVCC $FLAGS "$@" -c -o "$OBJ/${f%.*}.o" "$f" &
    pids="$pids $!"
done
for p in $pids; do wait "$p"; done
objs=""
for f in $SRCS; do objs="$objs $OBJ/${f%.*}.o"; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o "${OUT:-../libmagical_b200.so}" $objs
