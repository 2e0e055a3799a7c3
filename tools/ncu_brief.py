"""Prints the handful of ncu raw metrics we track + stall samples per code segment. Usage: ncu_brief.py rep [--src]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg ", "sm__cycles_elapsed.avg.per_second", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread ", "dram__bytes_read.sum ", "dram__bytes_write.sum ", "lts__t_bytes.sum ",
        "sm__inst_executed_pipe_fmaheavy", "sm__inst_executed_pipe_fmalite", "smsp__inst_executed.sum ",
        "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]
for r in rows[2:]:
    print("==", r[hdr.index("Kernel Name")][:80] if "Kernel Name" in hdr else "")
    for h, u, v in zip(hdr, units, r):
        if any((h + " ").startswith(w) or h.startswith(w.strip()) and w.endswith("pct") for w in want):
            print(f"  {h} [{u}] = {v}")
if "--src" in sys.argv:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr = rows[1]; data = rows[2:]
    iS, iSrc, iE = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
    tot = sum(int(r[iS]) for r in data)
    print("total samples", tot)
    keys = ("LDTM", "STTM", "SYNCS", "UTCHMMA", "UTCBAR", "BAR.", "UTMA", "EXIT", "LDG", "STS", "LDS")
    acc = 0; start = 0
    for i, r in enumerate(data):
        s = int(r[iS]); t = r[iSrc].strip()
        if any(k in t for k in keys):
            if acc > tot * 0.003: print(f"   [{start}-{i-1}] block samples={acc} ({100*acc/tot:.1f}%)")
            if s > tot * 0.003 or "SYNCS.PHASECHK" in t: print(f"{i} samples={s} ({100*s/tot:.1f}%) exec={r[iE]} {t[:80]}")
            acc = 0; start = i + 1
        else:
            acc += s
            if "BRA" in t and s > tot * 0.003: print(f"{i} samples={s} ({100*s/tot:.1f}%) exec={r[iE]} {t[:80]}")
