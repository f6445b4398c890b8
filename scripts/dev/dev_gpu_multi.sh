#!/bin/bash
# usage: dev_gpu_multi.sh N
N=${1:-2}
mkdir -p gpurun_out
{
nvidia-smi topo -m 2>&1 | head -14
echo "=== pytest multigpu"; timeout 900 python -m pytest tests/test_multigpu.py -x -q 2>&1 | tail -25
echo "=== headline columnwise N=$N"
timeout 600 python -m ddlb_b200.launcher -np $N --timeout 500 scripts/run_benchmark.py scripts/config_headline_columnwise.json 2>&1 | grep -v "^Running\|^ *implementation\|^$" | tail -25
echo "=== headline rowwise N=$N"
timeout 600 python -m ddlb_b200.launcher -np $N --timeout 500 scripts/run_benchmark.py scripts/config_headline_rowwise.json 2>&1 | grep -v "^Running\|^ *implementation\|^$" | tail -15
cp results/*.csv gpurun_out/ 2>/dev/null
} 2>&1 | tee gpurun_out/multi_$N.log
