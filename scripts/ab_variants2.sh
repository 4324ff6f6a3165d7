#!/bin/bash
# A/B of variant libraries with an arbitrary knob: KNOB=br_regs VALUES=16,8 bash scripts/ab_variants2.sh
mkdir -p gpurun_out; : > gpurun_out/ab_variants.jsonl
for so in tfhe_b200/variants/*.so; do
  name=$(basename $so .so)
  LUX_TFHE_LIB=$PWD/$so timeout 300 python scripts/ab_knob.py ${KNOB:-br_regs} --values ${VALUES:-16,8} --rounds ${ROUNDS:-3} 2>gpurun_out/ab_$name.err | sed "s/^{/{\"variant\": \"$name\", /" >> gpurun_out/ab_variants.jsonl
done
cat gpurun_out/ab_variants.jsonl
