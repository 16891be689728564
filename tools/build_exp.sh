#!/bin/bash
# development aid: build several FAST (FP32 kinematic only) variants of the library in parallel, one per "-D..." argument,
# into build_exp/lib_<i>.so; run any of them with MPPI_B200_LIB=build_exp/lib_<i>.so (see _capi.py)
cd "$(dirname "$0")/../motionplanning_b200/csrc"
i=0
for defs in "$@"; do
  ( nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -ccbin /usr/bin/g++ -Xcompiler -fPIC,-Wall,-Wno-unused-function \
      --expt-relaxed-constexpr -Xptxas -v -DMPPI_DEV_FAST $defs -shared -o ../../build_exp/lib_$i.so mppi_api.cu -ldl > ../../build_exp/build_$i.log 2>&1; echo "variant $i ($defs): rc=$?" ) &
  i=$((i+1))
done
wait
