#!/bin/bash
# Round-2 1-GPU validation of the rebuilt collectives (emulated ranks), fp32 (3xTF32), the fused
# epilogue activation and the no-alloc entry point; then timings. Every step has its own timeout
# and runs in its own process so that a trap in one kernel cannot hide the other results.
mkdir -p gpurun_out
{
for f in tests/test_gemm_gpu.py tests/test_fused_emulated_gpu.py tests/test_fp32_gpu.py \
         tests/test_models_gpu.py tests/test_trace_gpu.py tests/test_harness_gpu.py tests/test_zz_no_alloc.py; do
  echo "=== $f"
  timeout 900 python -m pytest $f -q -x 2>&1 | tail -12
done
echo "=== perf: bf16 headline / small / fp32"
timeout 300 python scripts/dev/dev_gemm_check.py --cta-group 2 --block-n 256 --b-layout kn --shapes 4096x1024x1024,65536x1024x1024,8192x1024x8192 --dtypes bfloat16 --perf
timeout 300 python scripts/dev/dev_gemm_check.py --cta-group 2 --block-n 256 --b-layout nk --shapes 8192x8192x1024,16384x4096x4096 --dtypes float32 --perf
timeout 300 python scripts/dev/dev_gemm_check.py --cta-group 2 --block-n 256 --b-layout nk --shapes 8192x8192x1024 --dtypes float32 --perf --fp32-mode tf32
echo "=== bench N=1"
timeout 300 python bench.py --steps 100 --warmup 10 2>gpurun_out/bench.err | tee gpurun_out/bench_n1.json | cut -c1-600
tail -3 gpurun_out/bench.err
} 2>&1 | tee gpurun_out/r2_s1.log
