#!/bin/bash
# ncu captures (plain flagship kernel, one emulated AG_before launch, one emulated RS launch),
# exported to text on the box: the .ncu-rep files together exceed what gpurun copies back.
mkdir -p gpurun_out/ncu
cd gpurun_out/ncu
run() {  # name, kernel regex, skip, command...
  local name=$1 regex=$2 skip=$3; shift 3
  timeout 120 "$@" > ${name}_plain.log 2>&1 && \
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c 1 -o $name "$@" > ${name}_ncu.log 2>&1
  echo "$name rc=$?"
  if [ -f $name.ncu-rep ]; then
    ncu -i $name.ncu-rep --page raw --csv > ${name}_raw.csv 2>/dev/null
    ncu -i $name.ncu-rep --page details > ${name}_details.txt 2>/dev/null
    ncu -i $name.ncu-rep --page source --csv > ${name}_source.csv 2>/dev/null
    rm -f $name.ncu-rep
  fi
}
cd ../..
ROOT=$(pwd)
cd gpurun_out/ncu
run plain_bf16_65536x1024x1024 tp_gemm_kernel 0 python $ROOT/scripts/dev/dev_gemm_check.py --cta-group 2 --block-n 256 --b-layout kn --shapes 65536x1024x1024 --dtypes bfloat16
run emulated_ag_before_2ranks tp_gemm_emulated_kernel 2 python $ROOT/scripts/dev/ncu_emulated.py ag 2
run emulated_rs_2ranks tp_gemm_emulated_kernel 2 python $ROOT/scripts/dev/ncu_emulated.py rs 2
ls -la
