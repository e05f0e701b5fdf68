#!/bin/bash
# build variants of libb2ba.so that differ in -D switches of ba_window.cu:  tools/exp_build.sh name "-DWX_STASH=0 ..." ...
set -e
cd "$(dirname "$0")/.."
mkdir -p build_exp
python -m visualodometry_b200._build > /dev/null
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr"
OBJS=$(ls visualodometry_b200/build/*.o | grep -v ba_window.o)
while [ $# -gt 1 ]; do
  name=$1; defs=$2; shift 2
  ( nvcc $FLAGS $defs -Xptxas -v -c visualodometry_b200/csrc/ba_window.cu -o build_exp/ba_window_$name.o 2>&1 | grep -A2 "Li7ELi4ELi4ELi32ELi12ELi4ELi1EEELi0" | grep -E "spill|registers" | tr '\n' ' '; echo " <- $name";
    nvcc -shared -o build_exp/libb2ba_$name.so $OBJS build_exp/ba_window_$name.o -gencode arch=compute_100a,code=sm_100a ) &
done
wait
