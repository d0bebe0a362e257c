#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sparse_gpu.py tests/test_configs_gpu.py tests/test_host_layer_gpu.py tests/test_engine_golden_gpu.py -m gpu -x -q > gpurun_out/r2_pytest11.txt 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2_pytest11.txt
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --nq-sweep= > gpurun_out/r2_bench11_mma.json 2> gpurun_out/r2_bench11_mma.err; echo "bench mma rc=$?"
MSR_LIB_PATH=$PWD/mllm_sparse_retrieval_b200/libmsr_nomma.so python bench.py --steps 10 --warmup 3 --no-cpu-baseline --nq-sweep= > gpurun_out/r2_bench11_nomma.json 2> gpurun_out/r2_bench11_nomma.err; echo "bench nomma rc=$?"
python bench.py --n-docs 125000 --steps 10 --warmup 3 --no-cpu-baseline --nq-sweep= > gpurun_out/r2_bench11_mma_125k.json 2>> gpurun_out/r2_bench11_mma.err; echo "bench mma 125k rc=$?"
MSR_LIB_PATH=$PWD/mllm_sparse_retrieval_b200/libmsr_nomma.so python bench.py --n-docs 125000 --steps 10 --warmup 3 --no-cpu-baseline --nq-sweep= > gpurun_out/r2_bench11_nomma_125k.json 2>> gpurun_out/r2_bench11_nomma.err; echo "bench nomma 125k rc=$?"
