#!/bin/bash
# parity suite + bench (no profiler)
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 ${BENCH_ARGS:-} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench.json'))
    print({k:d[k] for k in ('value','ms_per_step','kernel_ms','gpu_launches')}, d['e2e']['value'], d['roofline_int']['frac'], d['roofline_int']['peak'], d['clocks'])
except Exception as ex: print("no bench json", ex)
PY
