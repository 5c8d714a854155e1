#!/bin/bash
# builds the development micro-benchmarks next to their sources (run from anywhere)
set -e
D=$(cd "$(dirname "$0")" && pwd)
NV="nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr"
$NV -DQ4K_TIMING -o $D/mma_ticks $D/mma_ticks.cu
$NV -DQ4K_TIMING -DQ4K_TEST_WAIT -o $D/mma_ticks_tw $D/mma_ticks.cu
$NV -o $D/quant_lat $D/quant_lat.cu
$NV -o $D/pdl_floor $D/pdl_floor.cu
$NV -o $D/imma_rate $D/imma_rate.cu
$NV -o $D/stream_bulk $D/stream_bulk.cu
$NV -o $D/umma_rate $D/umma_rate.cu
ls -la $D | grep -v "\.cu$\|\.sh$"
