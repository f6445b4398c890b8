#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out results
timeout 600 python -m pytest tests/test_fused_emulated_gpu.py -x -q -k allgather 2>&1 | tail -4
bash scripts/dev/dev_gpu_perf.sh $N
