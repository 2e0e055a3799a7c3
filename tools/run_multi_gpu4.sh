#!/bin/bash
# 4-GPU box: DDP training step of config 5 (block pair and module-level swap at 576x1024 bf16, module at 512x320 fp32).
OUT=gpurun_out/${1:-r2_multi4}
: > ${OUT}_train_ddp.jsonl
PORT=29911
for cfg in "576x1024 bf16" "576x1024 bf16 --module" "512x320 fp32 --module"; do
  set -- $cfg
  PORT=$((PORT+1))
  timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port $PORT tools/train_ddp_bench.py --res $1 --param-dtype $2 $3 2>> ${OUT}_train_ddp.err | grep '^{' >> ${OUT}_train_ddp.jsonl
done
python - <<PY
import json
for l in open("${OUT}_train_ddp.jsonl"):
    d = json.loads(l)
    print(d["config"], d["resolution"], d["param_dtype"], round(d["ms_step_ddp"], 2), round(d["ms_step_no_sync"], 2), round(d["ms_allreduce_alone"], 3), d["gradients_identical_on_all_ranks"], d["finite"], round(d["clip_frames_per_s"]))
PY
