#!/bin/bash
# development aid: build pbhc_b200/variant_<name>.so with extra -D flags for track.cu:  tools/variant.sh name -DF_TW=7 ...
name=$1; shift
d=pbhc_b200/build/var_$name; mkdir -p $d
for f in abi lookup fk gae motion_prep; do cp pbhc_b200/build/$f.cu.o $d/ 2>/dev/null; done
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v "$@" -c pbhc_b200/csrc/track.cu -o $d/track.cu.o 2>&1 | grep -A2 "track_kernelILb0" | grep "registers\|spill"
nvcc -shared -o pbhc_b200/variant_$name.so $d/*.o -cudart shared
