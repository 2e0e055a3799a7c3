#!/bin/bash
# targeted tests first (short timeout), then the whole GPU suite, smoke and the bench line -> gpurun_out/<tag>_*
tag=${1:-r2_x}; shift
if [ $# -gt 0 ]; then
  timeout 240 python -m pytest "$@" -x -q -m gpu 2>&1 | tail -5 || exit 1
fi
timeout 200 python tools/prof_train_ops.py 2>&1 | grep -E "layernorm|geglu" 
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4; timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3) > gpurun_out/${tag}_tests_smoke.txt
cat gpurun_out/${tag}_tests_smoke.txt
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
tail -c 400 gpurun_out/${tag}_bench.err
python - <<PY
import json
for l in open("gpurun_out/${tag}_bench.json"):
    if l.startswith("{"):
        d = json.loads(l)
        print({k: d[k] for k in ("value", "ms_per_step")}, d["e2e"]["value"], d["roofline"]["frac"], d["roofline"]["ms_per_launch"])
        t = d.get("train_step", {})
        print("train", t.get("ms_fwd_bwd"), t.get("module", {}).get("ms_fwd_bwd"), t.get("ms_by_entry_point"))
PY
