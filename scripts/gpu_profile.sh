#!/bin/bash
# ncu full capture of the blind rotate kernel on a short bench run (after the same command exited 0 without ncu)
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --batch ${PROF_BATCH:-1332}"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:blind_rotate -s 3 -c 1 -o gpurun_out/prof_br $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
tail -2 gpurun_out/ncu_full.log | cut -c1-300
