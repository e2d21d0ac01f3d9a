#!/bin/bash
# Debug build of the library with the attention phase counters (-DBQA_ATTN_TRACE) -> tools/_trace/libbqa_trace<suffix>.so
# usage: tools/build_trace.sh [suffix] [extra nvcc flags...]
set -e
cd "$(dirname "$0")/.."
mkdir -p tools/_trace
SUF=$1; shift || true
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr \
  -DBQA_ATTN_TRACE "$@" -shared -o tools/_trace/libbqa_trace$SUF.so ml-recipe-distributed-pytorch_b200/csrc/*.cu -lcuda
