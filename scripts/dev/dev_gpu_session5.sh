#!/bin/bash
mkdir -p gpurun_out
{
echo "=== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
echo "=== bench ours"; timeout 600 python bench.py --steps 200 --warmup 20 2>gpurun_out/bench_ours.err | tee gpurun_out/bench_ours_n1.json; tail -5 gpurun_out/bench_ours.err
echo "=== bench reference"; timeout 900 python bench.py --impl reference --steps 50 --warmup 5 2>gpurun_out/bench_ref.err | tee gpurun_out/bench_ref_n1.json; tail -15 gpurun_out/bench_ref.err
} 2>&1 | tee gpurun_out/session5.log
