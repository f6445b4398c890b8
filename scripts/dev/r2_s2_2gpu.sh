#!/bin/bash
# Round-2 two-GPU session: VMM/NVLS heap bring-up, real-NVLink correctness (poisoned validation,
# slow rank, changing inputs, barrier wrap), transport A/B, both bench arms. Every step has its
# own timeout; nothing here may wait on a rendezvous for long.
mkdir -p gpurun_out
export DDLB_STORE_TIMEOUT=120
{
echo "=== 1. VMM heap + multicast self-test + multimem transports"
timeout 500 python -m pytest tests/test_vmm_multigpu.py -x -q -s > gpurun_out/vmm_test.log 2>&1
vmm_rc=$?
tail -25 gpurun_out/vmm_test.log
if [ $vmm_rc -ne 0 ]; then export DDLB_SYMM_MODE=ipc; echo "VMM tests failed (rc=$vmm_rc): the rest runs with DDLB_SYMM_MODE=ipc"; fi
echo "=== 2. multi-GPU suite at world 2"
timeout 900 python -m pytest tests/test_multigpu.py -x -q -k "2" 2>&1 | tail -15
echo "=== 3. AG transports (comm only vs headline)"
timeout 300 python -m ddlb_b200.launcher -np 2 --timeout 280 scripts/dev/r2_ag_comm_only.py 2>&1 | grep -E "world|transport|Error|error" | head -40
echo "=== 4. RS transports"
timeout 300 python -m ddlb_b200.launcher -np 2 --timeout 280 scripts/dev/r2_rowwise_ab.py 2>&1 | grep -E "world|transport|Error|error" | head -40
echo "=== 5. bench ours N=2"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 50 --warmup 10 2>gpurun_out/bench_ours_2.err | tee gpurun_out/bench_ours_n2.json | cut -c1-3000
tail -5 gpurun_out/bench_ours_2.err
echo "=== 6. bench reference N=2"
DDLB_REFERENCE_TIMEOUT=300 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus 2 --steps 50 --warmup 10 2>gpurun_out/bench_ref_2.err | tee gpurun_out/bench_ref_n2.json | cut -c1-2000
grep -iE "error|timed out|Traceback" gpurun_out/bench_ref_2.err | head -5
} 2>&1 | tee gpurun_out/r2_s2.log
