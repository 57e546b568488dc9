#!/bin/sh
# Dev only: second copy of the library with per-phase clock64() timers in the full-pel
# search (-DB200_ME_TIMING), used as B200AV1_LIB=.../libb200av1_timing.so B200_ME_DEBUG=1.
set -e
cd "$(dirname "$0")/../aom-av1-lavish_b200"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
$NVCC $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -DB200_ME_TIMING \
  -c csrc/me.cu -o /tmp/me_timing.o
$NVCC $ARCH -shared -o libb200av1_timing.so csrc/txq.o csrc/txq_inst_?.o csrc/blockops.o \
  csrc/capi.o csrc/compact.o csrc/hadamard.o csrc/quant_tables.o /tmp/me_timing.o csrc/rtcd.o -lcudart
