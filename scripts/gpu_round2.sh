#!/bin/bash
# one GPU session: full parity suites, headline bench (+ reference arm), DRAM traffic of the dominant kernel at the
# bench batch (ncu, two metrics), C1 gates bench, C3-C5 radix ops/s, latency probe
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/pytest_gpu.log)"
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -2 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; echo "bench ref rc=$?"
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:blind_rotate_kernel -s 3 -c 1 -f -o gpurun_out/prof_traffic python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_traffic.log 2>&1; echo "ncu traffic rc=$?"
timeout 900 python scripts/bench_gates.py ${GATES_ARGS:-} > gpurun_out/bench_gates.log 2> gpurun_out/bench_gates.err; echo "gates rc=$?"; tail -2 gpurun_out/bench_gates.err
timeout 900 python scripts/bench_radix.py > gpurun_out/bench_radix.log 2> gpurun_out/bench_radix.err; echo "radix rc=$?"; tail -3 gpurun_out/bench_radix.err
timeout 300 python scripts/latency_probe.py > gpurun_out/latency_probe.txt 2>&1; echo "latency rc=$?"
cat gpurun_out/bench.json | cut -c1-1200
tail -1 gpurun_out/bench_gates.log | cut -c1-800
cat gpurun_out/bench_radix.log | cut -c1-400
