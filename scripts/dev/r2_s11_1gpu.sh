#!/bin/bash
mkdir -p gpurun_out
{
echo "=== two-pair clusters: numerics"
timeout 300 python -m pytest tests/test_gemm_gpu.py -q -x -k "two_pair" 2>&1 | tail -6
timeout 300 python -m pytest tests/test_fused_emulated_gpu.py -q -x -k "two_pair" 2>&1 | tail -6
echo "=== perf: one pair vs two pairs vs cuBLAS (burst)"
for P in 0 2; do
timeout 300 python scripts/dev/dev_gemm_check.py --cta-group 2 --block-n 256 --b-layout kn --cluster-pairs $P --shapes 65536x1024x1024,8192x1024x8192,8192x8192x8192,16384x16384x16384,16384x28672x8192 --dtypes bfloat16 --perf
done
} 2>&1 | tee gpurun_out/r2_s11.log
