#!/bin/bash
# 2-GPU box: DDP training step of config 5 (block pair and module-level swap) and bench.py at N = 2, plus the reference arm.
OUT=gpurun_out/${1:-r2_multi2}
: > ${OUT}_scale.jsonl
: > ${OUT}_train_ddp.jsonl
PORT=29711
for cfg in "576x1024 bf16" "576x1024 bf16 --module" "512x320 fp32 --module"; do
  set -- $cfg
  PORT=$((PORT+1))
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT tools/train_ddp_bench.py --res $1 --param-dtype $2 $3 2>> ${OUT}_train_ddp.err | grep '^{' >> ${OUT}_train_ddp.jsonl
done
PORT=$((PORT+1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus 2 --steps 20 --warmup 5 2>> ${OUT}_scale.err | grep '^{' >> ${OUT}_scale.jsonl
PORT=$((PORT+1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT bench.py --impl reference --gpus 2 --steps 3 --warmup 1 2>> ${OUT}_scale.err | grep '^{' >> ${OUT}_scale.jsonl
wc -l ${OUT}_scale.jsonl ${OUT}_train_ddp.jsonl
python - <<PY
import json
for f in ("${OUT}_train_ddp.jsonl", "${OUT}_scale.jsonl"):
    for l in open(f):
        d = json.loads(l)
        print({k: d[k] for k in d if k in ("impl", "n_gpus", "value", "ms_per_step", "step_ms", "config", "world_size", "ms_fwd_bwd", "allreduce_ms", "grads_identical", "mode", "res")})
PY
