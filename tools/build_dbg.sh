#!/bin/bash
# debug build of the library with per-phase timing of the fused kernels (never shipped: loaded only through QB200_LIB)
cd "$(dirname "$0")/../quorum_b200" && /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared \
  -Xcompiler -fPIC -DQB_PHASE_TIMING -o libquorum_b200_dbg.so csrc/quorum_b200.cu
