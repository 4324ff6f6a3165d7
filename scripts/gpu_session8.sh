#!/bin/bash
# The multi-GPU session (gpurun --gpus N): headline bench under torchrun (weak and strong scaling), the same through the C ABI
# alone, and BASELINE configs C3-C5 sharded inside the library with every output decrypted.
#   gpurun --gpus 8 --timeout 1500 -- 'bash scripts/gpu_session8.sh 8'
set -u
N=${1:-8}
mkdir -p gpurun_out
run_bench() { # name, extra args
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 $2 > gpurun_out/$1.json 2> gpurun_out/$1.err
  echo "$1 rc=$?"; cut -c1-700 gpurun_out/$1.json
}
# STAGES (default: all) picks among weak strong abi configs
STAGES=${STAGES:-weak strong abi configs}
has() { case " $STAGES " in *" $1 "*) return 0;; esac; return 1; }
has weak && run_bench bench_n${N}_weak ""
has strong && run_bench bench_n${N}_strong "--scaling strong --c5 0"
has abi && timeout 600 python scripts/bench_multi_abi.py --gpus $N > gpurun_out/multi_abi_n${N}.json 2> gpurun_out/multi_abi_n${N}.err; echo "multi_abi rc=$?"; has abi && cut -c1-600 gpurun_out/multi_abi_n${N}.json
has configs && timeout 1500 python scripts/bench_configs.py --gpus $N ${CONFIG_ARGS:-} > gpurun_out/configs_n${N}.jsonl 2> gpurun_out/configs_n${N}.err; echo "configs rc=$?"; tail -3 gpurun_out/configs_n${N}.err; cut -c1-330 gpurun_out/configs_n${N}.jsonl
nvidia-smi --query-gpu=index,name,clocks.sm,power.draw --format=csv,noheader | head -8
