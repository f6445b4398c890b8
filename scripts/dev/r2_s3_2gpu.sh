#!/bin/bash
# Two-GPU session 3: VMM/NVLS heap (fd exchange fixed), transports after the relaxed-poll fix.
mkdir -p gpurun_out
export DDLB_STORE_TIMEOUT=120
{
echo "=== 1. VMM heap + multicast self-test + multimem transports"
timeout 500 python -m pytest tests/test_vmm_multigpu.py -x -q -s > gpurun_out/vmm_test.log 2>&1
vmm_rc=$?
tail -25 gpurun_out/vmm_test.log
if [ $vmm_rc -ne 0 ]; then export DDLB_SYMM_MODE=ipc; echo "VMM tests failed (rc=$vmm_rc): the rest runs with DDLB_SYMM_MODE=ipc"; fi
echo "=== 2. slow rank / changing inputs (all transports incl. multimem when available)"
timeout 300 python -m pytest tests/test_multigpu.py -x -q -k "slow_rank and 2" 2>&1 | tail -5
echo "=== 3. AG transports (comm only vs headline)"
timeout 300 python -m ddlb_b200.launcher -np 2 --timeout 280 scripts/dev/r2_ag_comm_only.py 2>&1 | grep -E "world|transport|Error|error" | head -40
echo "=== 4. RS transports"
DDLB_AB_LLAMA=0 timeout 300 python -m ddlb_b200.launcher -np 2 --timeout 280 scripts/dev/r2_rowwise_ab.py 2>&1 | grep -E "world|transport|Error|error" | head -40
echo "=== 5. bench ours N=2"
DDLB_BENCH_BASELINES=0 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 50 --warmup 10 2>gpurun_out/bench_ours_2.err | tee gpurun_out/bench_ours_n2.json | cut -c1-1500
tail -3 gpurun_out/bench_ours_2.err
python -c "from ddlb_b200.parallel.topology import max_active_clusters; print('max active clusters', max_active_clusters())"
} 2>&1 | tee gpurun_out/r2_s3.log
