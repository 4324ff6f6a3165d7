#!/bin/bash
# A/B of variant libraries on ONE box: for each tfhe_b200/variants/*.so run ab_knob (both gadget instantiations, 4 rounds)
mkdir -p gpurun_out; : > gpurun_out/ab_variants.jsonl
for so in tfhe_b200/variants/*.so; do
  name=$(basename $so .so)
  LUX_TFHE_LIB=$PWD/$so timeout 300 python scripts/ab_knob.py generic_gadget --rounds ${ROUNDS:-4} 2>gpurun_out/ab_$name.err | sed "s/^{/{\"variant\": \"$name\", /" >> gpurun_out/ab_variants.jsonl
done
cat gpurun_out/ab_variants.jsonl
