#!/bin/bash
# BN / GEGLU-block sweep of the forward GEMM at the L0 shapes (needs tools/micro/libtvi_tune.so)
cd "$(dirname "$0")/../.."
export PROF_LIB=tools/micro/libtvi_tune.so
for h in 128 64 32; do echo "== TVI_GEGLU_BLOCK=$h"; TVI_GEGLU_BLOCK=$h timeout 60 python tools/prof_ff.py ffin | tail -1; done
for bn in 256 192 160 128 96 64; do echo "== TVI_GEMM_BN=$bn"; TVI_GEMM_BN=$bn timeout 90 python tools/prof_ff.py ffin_plain,qkv,ffout,oproj | tail -4; done
