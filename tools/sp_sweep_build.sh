#!/bin/bash
# builds variant libraries of the streamed product (compile-time knobs) into tools/variants/ for tools/sp_sweep.py
cd "$(dirname "$0")/../visual_marker_mapping_b200/csrc"
mkdir -p ../../tools/variants
build() { # name flags...
  name=$1; shift
  nvcc -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -shared -o ../../tools/variants/libvmm_ba_$name.so ba_engine.cu -ldl > /dev/null 2>&1 || echo "FAILED $name"
}
build base &
build rel0 -DVMM_SP_RELEASE_MODE=0 &
build rel2 -DVMM_SP_RELEASE_MODE=2 &
build g3w4 -DVMM_SP_GROUPS=3 -DVMM_SP_GROUP_WARPS=4 &
wait
build g3w5 -DVMM_SP_GROUPS=3 -DVMM_SP_GROUP_WARPS=5 &
build g4w3 -DVMM_SP_GROUPS=4 -DVMM_SP_GROUP_WARPS=3 &
build g2w6 -DVMM_SP_GROUPS=2 -DVMM_SP_GROUP_WARPS=6 &
build ring24 -DVMM_SP_RING=24 &
wait
build ring16 -DVMM_SP_RING=16 &
build invv2 -DVMM_INV_SWEEP_V2=1 &
build invlook -DVMM_INV_LOOKAHEAD=1 &
build g3w4r24 -DVMM_SP_GROUPS=3 -DVMM_SP_GROUP_WARPS=4 -DVMM_SP_RING=24 &
wait
ls -la ../../tools/variants/
