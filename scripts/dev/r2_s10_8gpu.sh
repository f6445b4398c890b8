#!/bin/bash
# Eight-GPU session: correctness at world 8 on real NVLink, transports, traces, both bench arms,
# the BASELINE.json shape sweep and Llama shapes. Every step has its own timeout.
N=${1:-8}
mkdir -p gpurun_out
export DDLB_STORE_TIMEOUT=120
{
if [ "$N" = "2" ]; then
echo "=== 0. VMM heap + NVLS transports (2 ranks)"
timeout 400 python -m pytest tests/test_vmm_multigpu.py -x -q 2>&1 | tail -3
fi
echo "=== 1. multi-GPU suite at world $N"
if [ "${DDLB_SESSION_QUICK:-0}" = "1" ]; then
  timeout 300 python -m pytest tests/test_multigpu.py -x -q -k "(slow_rank or rowwise_fused) and $N" 2>&1 | tail -5
else
  timeout 600 python -m pytest tests/test_multigpu.py -x -q -k "$N" 2>&1 | tail -8
fi
echo "=== 2. bench ours N=$N"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 50 --warmup 10 2>gpurun_out/bench_ours_$N.err | tee gpurun_out/bench_ours_n$N.json | cut -c1-400
tail -3 gpurun_out/bench_ours_$N.err | cut -c1-300
echo "=== 3. bench reference N=$N"
DDLB_REFERENCE_TIMEOUT=300 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus $N --steps 50 --warmup 10 2>gpurun_out/bench_ref_$N.err | tee gpurun_out/bench_ref_n$N.json | cut -c1-400
grep -iE "error|timed out|Traceback" gpurun_out/bench_ref_$N.err | head -5
if [ "${DDLB_SESSION_NO_AB:-0}" != "1" ]; then
echo "=== 4. AG transports"
timeout 300 python -m ddlb_b200.launcher -np $N --timeout 280 scripts/dev/r2_ag_comm_only.py 2>&1 | grep -E "world|transport|Error|error" | cut -c1-170
echo "=== 5. RS transports"
timeout 300 python -m ddlb_b200.launcher -np $N --timeout 280 scripts/dev/r2_rowwise_ab.py 2>&1 | grep -E "world|transport|Error|error"
echo "=== 6. traces"
timeout 200 python -m ddlb_b200.launcher -np $N --timeout 180 scripts/dev/r2_trace_ag.py 2>&1 | grep -vE "^\[launcher\]"
fi
echo "=== 7. sweep + Llama shapes"
timeout 500 python -m ddlb_b200.launcher -np $N --timeout 480 scripts/sweep.py --llama --out gpurun_out/sweep_d$N.csv 2>&1 | grep -vE "^\[launcher\]" | tail -150
} 2>&1 | tee gpurun_out/r2_s10_$N.log
