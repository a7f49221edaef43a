#!/bin/bash
# build kernel variants (run in the container): tools/variants.sh "HB THREADS MINB" ...
mkdir -p /root/repo/gpurun_variants
cd /root/repo/graph-cut-ransac_b200/csrc
for v in "$@"; do set -- $v; name="hb$1_t$2_b$3";
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -diag-suppress 550 -shared -Xcompiler -fPIC -DGCR_HB=$1 -DGCR_THREADS=$2 -DGCR_MINB=$3 -Xptxas -v gcr_engine.cu -o /root/repo/gpurun_variants/lib_$name.so -lcudart 2>&1 | grep -A2 "k_score_fastILb0" | grep -E "Used|spill" | tr '\n' ' '; echo " <- $name"; done
