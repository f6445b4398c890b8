#!/bin/bash
# ncu capture of the flagship kernel on ONE GPU (run under gpurun). The plain run must exit 0
# first; ncu replays each kernel ~40x, so keep the program tiny. Never wrap a multi-rank run.
mkdir -p gpurun_out
CMD="python scripts/dev/dev_gemm_check.py --cta-group 2 --block-n 256 --b-layout kn --shapes 65536x1024x1024 --dtypes bfloat16"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tp_gemm_kernel -c 1 \
    -o gpurun_out/tp_gemm_g2_bn256 $CMD > gpurun_out/ncu.log 2>&1
tail -5 gpurun_out/ncu_plain.log gpurun_out/ncu.log
