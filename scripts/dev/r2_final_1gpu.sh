#!/bin/bash
# Final 1-GPU session: the whole GPU suite, the driver's bench line, ncu captures of the plain
# flagship kernel and of one emulated AG_before / RS launch (each only after the same command
# exited 0 without ncu).
mkdir -p gpurun_out
{
echo "=== pytest -m gpu"
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
echo "=== smoke"
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "=== bench N=1 (ours, reference)"
timeout 300 python bench.py --steps 100 --warmup 10 2>gpurun_out/bench_1.err | tee gpurun_out/bench_ours_n1.json | cut -c1-300
timeout 300 python bench.py --impl reference --steps 50 --warmup 5 2>gpurun_out/bench_ref_1.err | tee gpurun_out/bench_ref_n1.json | cut -c1-300
echo "=== perf"
timeout 300 python scripts/dev/dev_gemm_check.py --cta-group 2 --block-n 256 --b-layout kn --shapes 4096x1024x1024,65536x1024x1024,8192x1024x8192,8192x8192x8192 --dtypes bfloat16 --perf
timeout 300 python scripts/dev/dev_gemm_check.py --cta-group 2 --block-n 256 --b-layout nk --shapes 8192x8192x1024 --dtypes float32 --perf
echo "=== ncu"
CMD1="python scripts/dev/dev_gemm_check.py --cta-group 2 --block-n 256 --b-layout kn --shapes 65536x1024x1024 --dtypes bfloat16"
timeout 120 $CMD1 > gpurun_out/ncu_plain.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:tp_gemm_kernel -c 1 -o gpurun_out/ncu_plain_bf16 $CMD1 > gpurun_out/ncu1.log 2>&1; echo "plain rc=$?"
for MODE in ag rs; do
  CMD="python scripts/dev/ncu_emulated.py $MODE 2"
  timeout 120 $CMD > gpurun_out/ncu_${MODE}_plain.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:tp_gemm_emulated_kernel -s 2 -c 1 -o gpurun_out/ncu_emulated_$MODE $CMD > gpurun_out/ncu_$MODE.log 2>&1; echo "$MODE rc=$?"
done
} 2>&1 | tee gpurun_out/r2_final_1gpu.log
