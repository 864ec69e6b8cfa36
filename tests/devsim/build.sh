#!/bin/bash
# tests/devsim/build.sh — builds the CPU test double tests/devsim/libcfdnn_devsim.so
# (TEST INFRASTRUCTURE: the csrc sources compiled as plain C++ against tests/devsim/cuda_runtime.h).
set -e
here="$(cd "$(dirname "$0")" && pwd)"
root="$(cd "$here/../.." && pwd)"
src="$root/cfd-nn_b200/csrc"
out="$here/libcfdnn_devsim.so"
objs=()
mkdir -p "$here/obj"
for f in "$src"/*.cu; do
  o="$here/obj/$(basename "$f" .cu).o"
  if [ ! -f "$o" ] || [ -n "$(find "$src" "$here"/*.h "$here"/*.cpp "$root/include" -newer "$o" -print -quit)" ]; then
    g++ -std=c++17 -O2 -g -fPIC -ffp-contract=off -DCFDNN_DEVSIM $EXTRA -I"$here" -I"$root/include" -x c++ -c "$f" -o "$o" &
  fi
  objs+=("$o")
done
wait
g++ -std=c++17 -O2 -g -fPIC -DCFDNN_DEVSIM -I"$here" -c "$here/devsim.cpp" -o "$here/obj/devsim.o"
g++ -shared -o "$out" "${objs[@]}" "$here/obj/devsim.o"
echo "built $out"
