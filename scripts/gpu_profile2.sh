#!/bin/bash
# parity (quick) + launch list + full ncu capture (with source) of the dominant kernel
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; echo "parity rc=$? $(tail -1 gpurun_out/pytest_parity.log)"
timeout 300 python scripts/latency_probe.py > gpurun_out/latency_probe.txt 2>&1; grep "B=   1\|B= 148" gpurun_out/latency_probe.txt
CMD="python bench.py --steps 2 --warmup 3 --no-cpu"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
CMD2="python bench.py --steps 2 --warmup 3 --no-cpu --batch 1332"
timeout 600 $CMD2 > gpurun_out/plain2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:blind_rotate_kernel -s 3 -c 1 -o gpurun_out/prof_br $CMD2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out/*.ncu-rep
