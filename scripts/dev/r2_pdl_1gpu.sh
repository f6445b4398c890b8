#!/bin/bash
# One-GPU check of programmatic dependent launch: GEMM tests, on/off timing, bench N=1.
mkdir -p gpurun_out
{
echo "=== 1. GEMM + model tests"
timeout 300 python -m pytest tests/test_gemm_gpu.py tests/test_models_gpu.py tests/test_harness_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -4
echo "=== 2. PDL on/off, back-to-back launches"
timeout 200 python scripts/dev/r2_pdl_ab.py 2>&1 | tail -12
echo "=== 3. bench N=1"
timeout 300 python bench.py 2>gpurun_out/pdl_bench.err | tee gpurun_out/pdl_bench_n1.json | cut -c1-300
tail -2 gpurun_out/pdl_bench.err | cut -c1-200
echo "=== 4. bench N=1 with DDLB_PDL=0"
DDLB_PDL=0 DDLB_BENCH_BASELINES=0 timeout 300 python bench.py 2>/dev/null | tee gpurun_out/nopdl_bench_n1.json | cut -c1-300
} 2>&1 | tee gpurun_out/r2_pdl_1gpu.log
