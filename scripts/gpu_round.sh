#!/bin/bash
# one GPU session: full parity suites, headline bench, C1 gates bench + trials, C3-C5 radix ops/s
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/pytest_gpu.log)"
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -2 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; echo "bench ref rc=$?"
timeout 900 python scripts/bench_gates.py ${GATES_ARGS:-} > gpurun_out/bench_gates.log 2> gpurun_out/bench_gates.err; echo "gates rc=$?"; tail -2 gpurun_out/bench_gates.err
timeout 900 python scripts/bench_radix.py > gpurun_out/bench_radix.log 2> gpurun_out/bench_radix.err; echo "radix rc=$?"; tail -3 gpurun_out/bench_radix.err
cat gpurun_out/bench.json | cut -c1-1500
tail -1 gpurun_out/bench_gates.log | cut -c1-800
cat gpurun_out/bench_radix.log | cut -c1-260
