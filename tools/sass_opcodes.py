#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove (or disprove) a Blackwell-native kernel, from
`cuobjdump -sass libtvi.so` (B200_PROFILING.md "What proves a Blackwell-native kernel").  Writes profiles/sass_opcodes.md.

    python tools/sass_opcodes.py [path/to/libtvi.so]
"""
import re
import subprocess
import sys
from collections import Counter, OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
so = Path(sys.argv[1]) if len(sys.argv) > 1 else ROOT / "timelapse_video_interpolation_b200" / "libtvi.so"
COLS = ["UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "HMMA", "LDGSTS", "MUFU"]
txt = subprocess.run(["cuobjdump", "-sass", str(so)], capture_output=True, text=True, check=True).stdout
kernels = OrderedDict()
cur = None
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = name.replace("void ", "").replace("tvi::(anonymous namespace)::", "").replace("(anonymous namespace)::", "")
        name = re.sub(r"\(CUtensorMap_st.*|\((?:tvi::|__nv_bfloat16|float|unsigned|int|long|char|const).*", "", name)
        cur = kernels.setdefault(name, Counter())
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_]+)", line)
    if m and cur is not None:
        op = m.group(1)
        cur["_total"] += 1
        for c in COLS:
            if op.startswith(c):
                cur[c] += 1
out = ["# SASS opcode summary of libtvi.so (sm_100a)", "",
       "`python tools/sass_opcodes.py` — counts of the mnemonics B200_PROFILING.md lists: `UTCHMMA` = tcgen05.mma, `LDTM` / `STTM` =",
       "tcgen05.ld / st, `UTMALDG` / `UTMASTG` / `UBLKCP` = TMA (bulk tensor / bulk) copies, `HMMA` = legacy mma.sync, `LDGSTS` = cp.async.",
       "Kernels without any tensor-core instruction are the streaming (HBM-bound) kernels.", "",
       "| kernel | instr | " + " | ".join(COLS) + " | class |", "|---|---:|" + "---:|" * len(COLS) + "---|"]
for name, c in sorted(kernels.items(), key=lambda kv: (-kv[1]["UTCHMMA"], -kv[1]["HMMA"], kv[0])):
    cls = "tcgen05 / TMEM / TMA" if c["UTCHMMA"] else ("mma.sync" if c["HMMA"] else "streaming")
    out.append(f"| `{name}` | {c['_total']} | " + " | ".join(str(c[k]) if c[k] else "" for k in COLS) + f" | {cls} |")
(ROOT / "profiles" / "sass_opcodes.md").write_text("\n".join(out) + "\n")
print("\n".join(out[:8]))
print(f"... {len(kernels)} kernels -> profiles/sass_opcodes.md")
