#!/bin/sh
# Dev only: second copy of the library with per-phase clock64() timers in the full-pel
# search (-DB200_ME_TIMING), read by tools/me_phase_timing.py (B200AV1_LIB=.../libb200av1_timing.so).
set -e
cd "$(dirname "$0")/../aom-av1-lavish_b200"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
$NVCC $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -DB200_ME_TIMING $EXTRA \
  -c csrc/me.cu -o /tmp/me_timing.o
OBJS=$(ls csrc/*.o | grep -v '/me\.o$')
$NVCC $ARCH -shared -o ${OUT:-libb200av1_timing.so} $OBJS /tmp/me_timing.o -lcudart
