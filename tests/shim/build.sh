#!/bin/bash
# tests/shim/build.sh — builds the shim parity program twice (TEST INFRASTRUCTURE):
#   _build/shim_parity_sim : linked with the CPU test double  (-m "not gpu")
#   _build/shim_parity_gpu : linked with libcfdnn_b200.so      (-m gpu; travels to the GPU box as a built file)
# Needs the reference headers (/root/reference/include) and oracle/_ref/libnn_cfd_core.a.
set -e
here="$(cd "$(dirname "$0")" && pwd)"
root="$(cd "$here/../.." && pwd)"
REF="${REF:-/root/reference}"
mkdir -p "$here/_build"
common="-std=c++17 -O2 -w -I$REF/include -I$root/include $here/shim_parity.cpp $root/oracle/_ref/libnn_cfd_core.a -static-libstdc++ -static-libgcc"
g++ $common "$root/tests/devsim/libcfdnn_devsim.so" -Wl,-rpath,'$ORIGIN/../../devsim' -o "$here/_build/shim_parity_sim"
g++ $common "$root/cfd-nn_b200/libcfdnn_b200.so" -Wl,-rpath,'$ORIGIN/../../../cfd-nn_b200' -Wl,--allow-shlib-undefined -o "$here/_build/shim_parity_gpu"
echo "built $here/_build/shim_parity_{sim,gpu}"
