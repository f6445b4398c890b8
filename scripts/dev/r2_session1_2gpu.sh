#!/bin/bash
# Round-2 second GPU call (TWO GPUs, ~3 min = 6 GPU-min): reference arm under torchrun (the
# fix for the round-1 hang), both bench arms at N=2, rowwise trace.
mkdir -p gpurun_out
{
echo "=== bench reference N=2 (must finish in < 2 min)"
DDLB_REFERENCE_TIMEOUT=150 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus 2 --steps 30 --warmup 5 2>gpurun_out/bench_ref_2.err | tee gpurun_out/bench_ref_n2.json | cut -c1-400
grep -iE "error|timed out|Traceback" gpurun_out/bench_ref_2.err | head -5
echo "=== bench ours N=2"
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 100 --warmup 10 2>gpurun_out/bench_ours_2.err | tee gpurun_out/bench_ours_n2.json | cut -c1-400
echo "=== rowwise variants A/B (after they passed in emulation)"
timeout 200 python -m ddlb_b200.launcher -np 2 --timeout 180 scripts/dev/r2_rowwise_ab.py 2>&1 | grep variant
echo "=== VMM heap + multicast self-test"
DDLB_RUN_EXPERIMENTAL=1 timeout 200 python -m pytest tests/test_vmm_experimental_gpu.py -q -s 2>&1 | tail -12
echo "=== multigpu tests"
timeout 300 python -m pytest tests/test_multigpu.py -x -q -k "2" 2>&1 | tail -4
} 2>&1 | tee gpurun_out/r2_session1.log
