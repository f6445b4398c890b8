#!/bin/bash
mkdir -p gpurun_out
{
for mode in plain ag_after rs ag_before; do for g in 1 2; do
  echo "=== $mode g=$g"; timeout 90 python scripts/dev/dev_emu_check.py --mode $mode --cta-group $g 2>&1 | tail -12
done; done
echo "=== sanitizer (ag_before g=1)"
timeout 300 compute-sanitizer --tool memcheck --print-limit 5 python scripts/dev/dev_emu_check.py --mode ag_before --cta-group 1 --epochs 1 2>&1 | grep -v "^$" | tail -60
} 2>&1 | tee gpurun_out/emu_check.log
