#!/bin/bash
# Builds the stand-alone microbenchmarks / debug harnesses under tools/micro (binaries are git-ignored).
set -e
cd "$(dirname "$0")"
CS=../../timelapse_video_interpolation_b200/csrc
ARCH="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr"
nvcc $ARCH -o pipe_rates pipe_rates.cu
nvcc $ARCH -o softmax_step_bench softmax_step_bench.cu
nvcc $ARCH -o tmem_bw tmem_bw.cu
SRC="-DTVI_TC_POLY_SWEEP tc_timeline.cu $CS/attn_spatial_tc.cu $CS/attn_spatial.cu $CS/tvi_host.cu"
nvcc $ARCH -DTVI_TC_TIMELINE -o tc_timeline $SRC
nvcc $ARCH -o tc_time $SRC
nvcc $ARCH -DTVI_DEBUG_WAIT -o tc_time_dbg $SRC
echo built
