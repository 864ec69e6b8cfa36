#!/bin/bash
# cfd-nn_b200/build.sh — builds libcfdnn_b200.so for sm_100a (in-tree; travels to the GPU box).
# -fmad=false: the FP64 stencils must track the reference's x86 CPU build (no FMA contraction,
# reference CMakeLists.txt:221-223) to rounding; the kernels are HBM-bound so it costs nothing.
set -e
here="$(cd "$(dirname "$0")" && pwd)"
src="$here/csrc"
out="$here/libcfdnn_b200.so"
mkdir -p "$here/build"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC -I$here/../include ${EXTRA_NVCC_FLAGS}"
pids=()
objs=()
for f in "$src"/*.cu; do
  o="$here/build/$(basename "$f" .cu).o"
  objs+=("$o")
  if [ ! -f "$o" ] || [ -n "$(find "$src" "$here/../include" -newer "$o" -print -quit)" ]; then
    $NVCC $FLAGS -c "$f" -o "$o" &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait "$p"; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o "$out" "${objs[@]}" -ldl
echo "built $out"
