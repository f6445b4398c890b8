#!/bin/bash
# One-GPU check of the wide-tile kernels: tests, then the A/B against 256x256 tiles and cuBLAS.
mkdir -p gpurun_out
{
echo "=== 1. wide-tile tests"
timeout 200 python -m pytest tests/test_gemm_gpu.py tests/test_fused_emulated_gpu.py -x -q -k "wide or auto_tile or dependent" 2>&1 | tail -6
echo "=== 2. wide vs narrow vs cuBLAS"
timeout 200 python scripts/dev/r2_wide_ab.py 2>&1 | tail -14
} 2>&1 | tee gpurun_out/r2_wide_1gpu.log
