#!/bin/bash
# One multi-GPU box: bench.py at N = 1, 2, 4, ... (weak scaling incl. e2e and its host ceiling) and the DDP training step
# of config 5 at both resolutions.  Usage: tools/run_multi_gpu.sh <max N> <out prefix>
set -u
NMAX=${1:-2}
OUT=${2:-gpurun_out/r2_multi}
: > ${OUT}_scale.jsonl
: > ${OUT}_train_ddp.jsonl
PORT=29511
for N in 1 2 4 8; do
  [ $N -gt $NMAX ] && break
  PORT=$((PORT+1))
  if [ $N -eq 1 ]; then
    timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-extras --no-cpu-baseline >> ${OUT}_scale.jsonl 2>> ${OUT}_scale.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 20 --warmup 5 >> ${OUT}_scale.jsonl 2>> ${OUT}_scale.err
  fi
  for cfg in "576x1024 bf16" "512x320 fp32"; do
    set -- $cfg
    PORT=$((PORT+1))
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT tools/train_ddp_bench.py --res $1 --param-dtype $2 >> ${OUT}_train_ddp.jsonl 2>> ${OUT}_train_ddp.err
  done
done
nvidia-smi topo -m > ${OUT}_topo.txt 2>&1
wc -l ${OUT}_scale.jsonl ${OUT}_train_ddp.jsonl
tail -3 ${OUT}_train_ddp.err
