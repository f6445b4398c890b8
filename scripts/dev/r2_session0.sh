#!/bin/bash
# Round-2 first GPU call (ONE GPU, ~3 min): validate everything written without a GPU at the
# end of round 1, then capture the flagship kernel with ncu. Every step has its own timeout.
mkdir -p gpurun_out
{
echo "=== 1. regular GPU suite"
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
echo "=== 2. experimental: TF32, trace, rowwise variants (each file on its own: a failure in one must not hide the others)"
for f in tests/test_tf32_experimental_gpu.py tests/test_trace_experimental_gpu.py tests/test_variants_experimental_gpu.py; do
  DDLB_RUN_EXPERIMENTAL=1 timeout 400 python -m pytest $f -q -s 2>&1 | tail -25
done
echo "=== 3. bench N=1 (serial e2e) and pipelined e2e"
timeout 300 python bench.py --steps 200 --warmup 20 2>gpurun_out/bench.err | tee gpurun_out/bench_n1.json | cut -c1-400
DDLB_BENCH_PIPELINED_E2E=1 timeout 300 python bench.py --steps 100 --warmup 10 2>gpurun_out/bench_pipe.err | tee gpurun_out/bench_n1_pipelined.json | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['e2e'])"
tail -3 gpurun_out/bench_pipe.err
echo "=== 4. ncu (flagship kernel, single launch)"
bash scripts/dev/dev_ncu.sh
} 2>&1 | tee gpurun_out/r2_session0.log
