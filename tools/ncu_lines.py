#!/usr/bin/env python
"""Per CUDA source line: stall samples, warp instructions, threads per instruction, from an
.ncu-rep captured with --import-source on.   usage: ncu_lines.py report.ncu-rep [top-n]"""
import csv, subprocess, sys, os
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 45
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows, hdr, fname = [], None, ""
for rec in csv.reader(out.splitlines()):
    if not rec:
        continue
    if rec[0] == "File Path":
        fname = os.path.basename(rec[1]); hdr = None; continue
    if rec[0] == "Function Name":
        continue
    if rec[0] == "Line No":
        hdr = rec; continue
    if hdr and rec[0] not in ("", "-"):
        d = {}
        for k, v in zip(hdr, rec):
            d.setdefault(k, v)          # the first "Source" column is the CUDA line
        d["file"] = fname
        rows.append(d)
key = lambda r, c: int(float(r.get(c) or 0)) if (r.get(c) or "0").replace(".", "").isdigit() else 0
tot = sum(key(r, "# Samples") for r in rows)
inst = sum(key(r, "Instructions Executed") for r in rows)
thr = sum(key(r, "Thread Instructions Executed") for r in rows)
print(f"{rep}: {tot} samples, {inst} warp instr, {thr / max(inst, 1):.1f} threads/instr")
stall_cols = [c for c in rows[0] if c.startswith("stall_") and "Not Issued" not in c]
mix = {c: sum(key(r, c) for r in rows) for c in stall_cols}
print("stall mix:", ", ".join(f"{k[6:]} {v * 100 // max(tot, 1)}%" for k, v in sorted(mix.items(), key=lambda kv: -kv[1])[:8]))
for r in sorted(rows, key=lambda r: -key(r, "# Samples"))[:top]:
    s = key(r, "# Samples"); ie = key(r, "Instructions Executed"); te = key(r, "Thread Instructions Executed")
    why = sorted(((key(r, c), c[6:]) for c in stall_cols), reverse=True)[:2]
    print(f"{r['file'][:18]:18s}:{r['Line No']:>4} smp {s * 100.0 / max(tot, 1):5.1f}% inst {ie * 100.0 / max(inst, 1):5.1f}% thr {te / max(ie, 1):4.1f}  {r['Source'].strip()[:88]:88s} {why[0][1]}:{why[0][0] * 100 // max(s, 1)}% {why[1][1]}:{why[1][0] * 100 // max(s, 1)}%")
