#!/bin/bash
mkdir -p gpurun_out
{
for g in 1 2; do for w in 2 8; do
  echo "=== ag_before g=$g w=$w"; timeout 90 python scripts/dev/dev_emu_check.py --mode ag_before --cta-group $g --world $w --epochs 3 2>&1 | tail -30
done; done
} 2>&1 | tee gpurun_out/emu_check2.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -30 | tee gpurun_out/pytest_gpu.log
