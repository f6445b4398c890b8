#!/bin/bash
# Release check on one GPU: whole GPU suite, smoke, both bench arms with default flags.
mkdir -p gpurun_out
{
echo "=== pytest -m gpu"
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
echo "=== smoke"
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
echo "=== bench N=1 (reference, ours)"
timeout 300 python bench.py --impl reference 2>gpurun_out/release_ref_1.err | tee gpurun_out/release_ref_n1.json | cut -c1-260
timeout 300 python bench.py 2>gpurun_out/release_ours_1.err | tee gpurun_out/release_ours_n1.json | cut -c1-260
} 2>&1 | tee gpurun_out/r2_release_1gpu.log
