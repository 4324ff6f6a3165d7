#!/bin/bash
# launch list of the bench command + full ncu capture (with source) of the dominant kernel at 3 waves
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
CMD2="python bench.py --steps 2 --warmup 3 --no-cpu --batch 888"
timeout 600 $CMD2 > gpurun_out/plain2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:blind_rotate_kernel -s 3 -c 1 -f -o gpurun_out/prof_br $CMD2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:lwe_keyswitch_mma -s 3 -c 1 -f -o gpurun_out/prof_ks python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_ks.log 2>&1
echo "ncu ks rc=$?"; ls -la gpurun_out/*.ncu-rep
