#!/bin/bash
# Final two-GPU confirmation on the release commit: both bench arms under torchrun, then the
# quick part of the multi-GPU suite. Every step has its own timeout.
mkdir -p gpurun_out
export DDLB_STORE_TIMEOUT=120
{
echo "=== 1. bench reference N=2"
DDLB_REFERENCE_TIMEOUT=150 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus 2 --steps 50 --warmup 10 2>gpurun_out/final_ref_2.err | tee gpurun_out/final_ref_n2.json | cut -c1-400
grep -iE "error|timed out|Traceback" gpurun_out/final_ref_2.err | head -5
echo "=== 2. bench ours N=2"
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 50 --warmup 10 2>gpurun_out/final_ours_2.err | tee gpurun_out/final_ours_n2.json | cut -c1-600
tail -3 gpurun_out/final_ours_2.err | cut -c1-300
echo "=== 3. multi-GPU suite, world 2, quick subset"
timeout 150 python -m pytest tests/test_multigpu.py -x -q -k "(slow_rank or rowwise_fused or tp_mlp) and 2" 2>&1 | tail -4
} 2>&1 | tee gpurun_out/r2_final_2gpu.log
