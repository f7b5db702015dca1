#!/bin/bash
# lab builds of libpcb200 with extra macros into build/abl/libpcb200_<name>.so (parallel):
#   scripts/abl_build.sh name1:"-DV9_RING=0" name2:"-DPCB200_ABL=3" ...
cd "$(dirname "$0")/../pencil_code_b200/csrc" || exit 1
mkdir -p ../../build/abl
for spec in "$@"; do
  name=${spec%%:*}; flags=${spec#*:}
  ( nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wno-attribute-alias $flags -shared \
      -o ../../build/abl/libpcb200_$name.so pcb200.cu pcb200_driver.cpp pcb200_snapshot.cpp -ldl > ../../build/abl/build_$name.log 2>&1; echo "built $name rc=$?" ) &
done
wait
