#!/bin/bash
# builds scratch/libmsr_<name>.so with extra -D flags for sparse.cu (kernel tuning experiments)
# usage: scratch/build_variant.sh name -DMSR_SPARSE_THREADS=256 -DMSR_SPARSE_MIN_CTAS=4
set -e
name=$1; shift
P=mllm_sparse_retrieval_b200
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -Xptxas -v -c $P/csrc/sparse.cu -o /tmp/sparse_$name.o 2>&1 | grep -A2 "sparse_score_topk" | tail -2
objs=$(ls $P/build/*.o | grep -v sparse.o)
nvcc -shared -o scratch/libmsr_$name.so /tmp/sparse_$name.o $objs -gencode arch=compute_100a,code=sm_100a -lcudart
echo built scratch/libmsr_$name.so
