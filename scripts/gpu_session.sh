#!/bin/bash
# One GPU session on the gpurun box: a list of stages, each writing under gpurun_out/.  The one runner for every measurement
# of this repo (it replaces the per-session one-off scripts of round 1).
#   gpurun --timeout 1500 -- 'bash scripts/gpu_session.sh tests_torus bench ncu_fft'
# stages:
#   tests          full -m gpu suite                 tests_torus   torus + torus-radix suites only
#   bench          headline bench + reference arm    bench_gold    the Goldilocks arithmetic on the same shape
#   quick          scripts/quick_torus.py $QUICK_ARGS (device-resident kernel timing; teams/variant sweeps)
#   ncu_fft        ncu --set full (+source) of the torus blind rotation at B=$NCU_B (default 4440) + launch list of the bench
#   launches       ncu launch list (gpu__time_duration) of the bench command
#   traffic        DRAM bytes of the dominant kernel at the bench batch
#   gate           scripts/decrypt_trials.py (>= 10^6 trials), torus then goldilocks
#   radix          scripts/bench_radix.py (C3-C5 op rates, single process)
#   latency        scripts/latency_probe.py
#   ubench         scripts/ubench/ubench_arith.cu
#   ncu_ks         ncu --set full of the tcgen05 key switch at B=$NCU_B
#   ks             key switch parity tests + scripts/ks_probe.py (mma.sync vs tcgen05 form)
set -u
mkdir -p gpurun_out
NCU_B=${NCU_B:-4440}
for stage in "$@"; do
  echo "=== stage $stage"
  case $stage in
    tests) timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/pytest_gpu.log)";;
    tests_torus) timeout 1500 python -m pytest tests/test_gpu_torus.py tests/test_gpu_torus_radix.py -m gpu -x -q > gpurun_out/pytest_torus.log 2>&1; echo "pytest rc=$? $(tail -3 gpurun_out/pytest_torus.log)";;
    bench) timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err; cut -c1-1500 gpurun_out/bench.json
           timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; echo "bench ref rc=$?";;
    bench_gold) timeout 900 python bench.py --arith goldilocks --steps 5 --warmup 3 > gpurun_out/bench_gold.json 2> gpurun_out/bench_gold.err; echo "bench gold rc=$?"; cut -c1-600 gpurun_out/bench_gold.json;;
    quick) timeout 600 python scripts/quick_torus.py ${QUICK_ARGS:-4440 0} 2>&1 | tee gpurun_out/quick.txt;;
    ncu_fft)
      timeout 900 ncu --set full --import-source on --clock-control none -k regex:blind_rotate_fft -s 2 -c 1 -f -o gpurun_out/fft_full python scripts/quick_torus.py $NCU_B 0 > gpurun_out/ncu_fft.log 2>&1; echo "ncu full rc=$?"
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --c5 0 > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?";;
    launches) timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --c5 0 > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?";;
    traffic) timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:blind_rotate_fft -s 3 -c 1 -f -o gpurun_out/prof_traffic python bench.py --steps 2 --warmup 3 --no-cpu --c5 0 > gpurun_out/ncu_traffic.log 2>&1; echo "ncu traffic rc=$?";;
    gate) timeout 1200 python scripts/decrypt_trials.py --arith torus --out gpurun_out/decrypt_trials_torus.json > gpurun_out/gate_torus.log 2>&1; echo "gate torus rc=$?"; tail -1 gpurun_out/gate_torus.log | cut -c1-600
          timeout 1200 python scripts/decrypt_trials.py --arith goldilocks --out gpurun_out/decrypt_trials_goldilocks.json > gpurun_out/gate_gold.log 2>&1; echo "gate goldilocks rc=$?";;
    radix) timeout 900 python scripts/bench_radix.py ${RADIX_ARGS:-} > gpurun_out/bench_radix.jsonl 2> gpurun_out/bench_radix.err; echo "radix rc=$?"; tail -3 gpurun_out/bench_radix.err; cut -c1-300 gpurun_out/bench_radix.jsonl;;
    ks) timeout 240 python -m pytest tests/test_gpu_torus.py tests/test_gpu_parity.py -m gpu -x -q -k "keyswitch" > gpurun_out/pytest_ks.log 2>&1; echo "pytest ks rc=$? $(tail -3 gpurun_out/pytest_ks.log)"
        timeout 240 python scripts/ks_probe.py torus > gpurun_out/ks_probe.txt 2>&1; echo "ks probe rc=$?"; tail -12 gpurun_out/ks_probe.txt;;
    ncu_ks) timeout 600 ncu --set full --import-source on --clock-control none -k regex:lwe_keyswitch_tc5 -s 2 -c 1 -f -o gpurun_out/ks_tc5_full python scripts/quick_torus.py $NCU_B 0 > gpurun_out/ncu_ks.log 2>&1; echo "ncu ks rc=$?";;
    latency) timeout 300 python scripts/latency_probe.py > gpurun_out/latency_probe.txt 2>&1; echo "latency rc=$?"; tail -8 gpurun_out/latency_probe.txt;;
    ubench) (cd scripts/ubench && nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o /tmp/ubench_arith ubench_arith.cu && /tmp/ubench_arith) > gpurun_out/arith_ubench.txt 2>&1; echo "ubench rc=$?";;
    *) echo "unknown stage $stage";;
  esac
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
