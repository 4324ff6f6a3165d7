#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/sweep.jsonl
for cfg in "4440 0" "2960 2" "1480 1" "1024 0" "16384 0" "65536 0"; do set -- $cfg
  timeout 900 python bench.py --steps 3 --warmup 3 --batch $1 --teams $2 --no-cpu >> gpurun_out/sweep.jsonl 2>> gpurun_out/sweep.err
done
python - <<'PY'
import json
for l in open('gpurun_out/sweep.jsonl'):
    d=json.loads(l); print(d['config']['batch_per_gpu'], d['config']['wave_ciphertexts'], round(d['value']), round(d['e2e']['value']), d['kernel_ms'], round(d['roofline_int']['frac'],3))
PY
