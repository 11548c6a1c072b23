#!/bin/bash
# builds variant libraries of the streamed product (compile-time knobs) into tools/variants/ for tools/sp_sweep.py
cd "$(dirname "$0")/../visual_marker_mapping_b200/csrc"
mkdir -p ../../tools/variants
build() { # name flags...
  name=$1; shift
  nvcc -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -shared -o ../../tools/variants/libvmm_ba_$name.so ba_engine.cu -ldl > /dev/null 2>&1 || echo "FAILED $name"
}
build base &
build g2w7r20 -DVMM_SP_RING=20 &
build g2w7r12 -DVMM_SP_RING=12 &
build g3w5r16 -DVMM_SP_GROUPS=3 -DVMM_SP_GROUP_WARPS=5 &
wait
build g3w5r20 -DVMM_SP_GROUPS=3 -DVMM_SP_GROUP_WARPS=5 -DVMM_SP_RING=20 &
build g3w5r12 -DVMM_SP_GROUPS=3 -DVMM_SP_GROUP_WARPS=5 -DVMM_SP_RING=12 &
build g3w5r16rel0 -DVMM_SP_GROUPS=3 -DVMM_SP_GROUP_WARPS=5 -DVMM_SP_RELEASE_MODE=0 &
build invv2 -DVMM_INV_SWEEP_V2=1 &
wait
build g2w8r16 -DVMM_SP_GROUPS=2 -DVMM_SP_GROUP_WARPS=8 &
build g4w4r16 -DVMM_SP_GROUPS=4 -DVMM_SP_GROUP_WARPS=4 &
wait
ls ../../tools/variants/ | tr '\n' ' '
