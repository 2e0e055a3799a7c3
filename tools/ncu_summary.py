#!/usr/bin/env python
"""Summarises an .ncu-rep (ncu --set full) and/or a launch-list CSV into small tracked files under profiles/.

    python tools/ncu_summary.py --rep gpurun_out/prof.ncu-rep --launches gpurun_out/launches.csv --out profiles/r1_block_L0
"""
import argparse
import csv
import io
import subprocess

WANT = [
    ("Kernel Name", "kernel"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs"), ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram_read"), ("dram__bytes_write.sum", "dram_write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_pct"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct"),
    ("smsp__inst_executed.sum", "inst"),
]


def rep_rows(path):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        d = {}
        for name, short in WANT:
            if name in hdr:
                i = hdr.index(name)
                d[short] = r[i] + (" " + units[i] if units[i] and short not in ("kernel",) else "")
        out.append(d)
    return out


def launch_rows(path):
    lines = [l for l in open(path).read().splitlines() if not l.startswith("==")]
    out = []
    for row in csv.DictReader(io.StringIO("\n".join(lines))):
        if row.get("Metric Name") == "gpu__time_duration.sum":
            out.append((row["ID"], row["Kernel Name"], float(row["Metric Value"].replace(",", "")), row["Metric Unit"]))
    return out


def short(k):
    k = k.replace("tvi::<unnamed>::", "").replace("void ", "")
    return k.split("(")[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rep")
    ap.add_argument("--launches")
    ap.add_argument("--out", required=True)
    ap.add_argument("--last", type=int, default=0, help="only the last N launches of the list (one step)")
    a = ap.parse_args()
    md = []
    if a.launches:
        rows = launch_rows(a.launches)
        if a.last:
            rows = rows[-a.last:]
        tot = sum(r[2] for r in rows)
        md.append(f"## launch list ({len(rows)} launches, total {tot / 1e6:.3f} ms; ncu gpu__time_duration, cold-cache, serialised)\n")
        md.append("| # | kernel | ns | share |\n|---|---|---:|---:|")
        for r in rows:
            md.append(f"| {r[0]} | {short(r[1])} | {r[2]:.0f} | {100 * r[2] / tot:.1f}% |")
        agg = {}
        for r in rows:
            agg.setdefault(short(r[1]), [0, 0.0])
            agg[short(r[1])][0] += 1
            agg[short(r[1])][1] += r[2]
        md.append("\n| kernel | launches | total ns | share |\n|---|---:|---:|---:|")
        for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            md.append(f"| {k} | {n} | {t:.0f} | {100 * t / tot:.1f}% |")
    if a.rep:
        rows = rep_rows(a.rep)
        md.append(f"\n## ncu --set full ({a.rep.split('/')[-1]})\n")
        cols = [s for _, s in WANT]
        md.append("| " + " | ".join(cols) + " |\n|" + "---|" * len(cols))
        for d in rows:
            d["kernel"] = short(d.get("kernel", ""))
            md.append("| " + " | ".join(d.get(c, "") for c in cols) + " |")
    open(a.out + ".md", "w").write("\n".join(md) + "\n")
    print("\n".join(md))


if __name__ == "__main__":
    main()
