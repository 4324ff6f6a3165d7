#!/bin/bash
# A/B the tuning variants built by scripts/build_variant.sh: parity on a subset, then bench each
set -u
mkdir -p gpurun_out; : > gpurun_out/variants.jsonl
for so in tfhe_b200/variants/*.so; do
  name=$(basename $so .so)
  LUX_TFHE_LIB=$PWD/$so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q  > gpurun_out/parity_$name.log 2>&1
  echo "$name parity rc=$? $(tail -1 gpurun_out/parity_$name.log)"
  LUX_TFHE_LIB=$PWD/$so timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu 2>> gpurun_out/variants.err | sed "s/^{/{\"variant\": \"$name\", /" >> gpurun_out/variants.jsonl
done
python - <<'PY'
import json
for l in open('gpurun_out/variants.jsonl'):
    d=json.loads(l); print(d['variant'], round(d['value']), round(d['e2e']['value']), d['kernel_ms'], round(d['roofline_int']['frac'],3), d['latency_single_pbs_ms'])
PY
