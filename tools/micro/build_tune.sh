#!/bin/bash
# libtvi with tuning hooks compiled in (-DTVI_GEMM_TUNE: TVI_GEGLU_BLOCK, TVI_GEMM_BN, TVI_LN_GENERIC) -> tools/micro/libtvi_tune<suffix>.so
# use: build_tune.sh [suffix] [extra nvcc flags...];  PROF_LIB=tools/micro/libtvi_tune.so TVI_GEGLU_BLOCK=64 python tools/prof_ff.py ffin
set -e
cd "$(dirname "$0")"
CS=../../timelapse_video_interpolation_b200/csrc
SUF="$1"; shift || true
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr -Xcompiler -fPIC -shared \
  -DTVI_GEMM_TUNE "$@" -o libtvi_tune$SUF.so $CS/*.cu -lcuda
