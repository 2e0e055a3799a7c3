#!/bin/bash
# builds tc_time from an alternative kernel source (A/B comparisons): build_ab.sh <kernel.cu> <out>
set -e
cd "$(dirname "$0")"
CS=../../timelapse_video_interpolation_b200/csrc
cp "$1" /tmp/ab_attn_spatial_tc.cu
sed -i 's#"tvi_common.cuh"#"'$(cd $CS && pwd)'/tvi_common.cuh"#' /tmp/ab_attn_spatial_tc.cu
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr -DTVI_TC_POLY_SWEEP -o "$2" tc_timeline.cu /tmp/ab_attn_spatial_tc.cu $CS/attn_spatial.cu $CS/tvi_host.cu
