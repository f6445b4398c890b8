#!/bin/bash
# ncu capture of the 3xTF32 kernel (fp32 8192x8192x1024, cta_group 2, BLOCK_N 256), exported to
# text on the box. Runs only after the same command exited 0 without ncu.
mkdir -p gpurun_out/ncu
ROOT=$(pwd)
cd gpurun_out/ncu
name=plain_fp32_3xtf32_8192x8192x1024
CMD="python $ROOT/scripts/dev/dev_gemm_check.py --cta-group 2 --block-n 256 --b-layout nk --shapes 8192x8192x1024 --dtypes float32"
timeout 100 $CMD > ${name}_plain.log 2>&1 && \
timeout 150 ncu --set full --clock-control none --import-source on -k regex:tp_gemm_kernel -c 1 -o $name $CMD > ${name}_ncu.log 2>&1
echo "$name rc=$?"
if [ -f $name.ncu-rep ]; then
  ncu -i $name.ncu-rep --page raw --csv > ${name}_raw.csv 2>/dev/null
  ncu -i $name.ncu-rep --page details > ${name}_details.txt 2>/dev/null
  ncu -i $name.ncu-rep --page source --csv > ${name}_source.csv 2>/dev/null
  rm -f $name.ncu-rep
fi
tail -3 ${name}_plain.log | cut -c1-300
ls -la
