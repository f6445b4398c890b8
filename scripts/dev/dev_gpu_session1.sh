#!/bin/bash
# First GPU session: topology, plain GEMM variants (each in its own process under timeout).
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -m ddlb_b200.parallel.topology > gpurun_out/topology.json 2>&1
tail -5 gpurun_out/topology.json
for g in 1 2; do for b in kn nk; do for bn in 256 128; do
  echo "=== cta_group=$g b=$b block_n=$bn"
  timeout 120 python scripts/dev/dev_gemm_check.py --cta-group $g --block-n $bn --b-layout $b --perf 2>&1 | tail -14
  echo "exit=$?"
done; done; done 2>&1 | tee gpurun_out/gemm_check.log
