#!/bin/bash
# 8-GPU box: DDP training step of config 5 at N = 4, 8 (both resolutions) and bench.py at N = 8 (e2e + host ceiling).
OUT=gpurun_out/r2_multi8
: > ${OUT}_scale.jsonl
: > ${OUT}_train_ddp.jsonl
PORT=29611
for N in 4 8; do
  for cfg in "576x1024 bf16" "512x320 fp32" "576x1024 bf16 --module" "512x320 fp32 --module"; do
    set -- $cfg
    PORT=$((PORT+1))
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT tools/train_ddp_bench.py --res $1 --param-dtype $2 $3 2>> ${OUT}_train_ddp.err | grep '^{' >> ${OUT}_train_ddp.jsonl
  done
done
for N in 4 8; do
PORT=$((PORT+1))
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 20 --warmup 5 2>> ${OUT}_scale.err | grep '^{' >> ${OUT}_scale.jsonl
done
nvidia-smi topo -m > ${OUT}_topo.txt 2>&1
wc -l ${OUT}_scale.jsonl ${OUT}_train_ddp.jsonl
