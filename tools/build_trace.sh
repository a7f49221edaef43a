#!/bin/bash
# Builds the diagnostic engine variant read by tools/trace_score.py: gpurun_variants/lib_trace.so = the engine with -DGCR_TRACE
# (per-CTA / per-warp %globaltimer stamps in the scoring kernels, gcr_debug_read_trace).  Not part of the product build.
set -e
cd "$(dirname "$0")/.."
mkdir -p gpurun_variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -diag-suppress 550 -shared -Xcompiler -fPIC \
     -DGCR_TRACE graph-cut-ransac_b200/csrc/gcr_engine.cu -o gpurun_variants/lib_trace.so -lcudart
echo built gpurun_variants/lib_trace.so
