#!/usr/bin/env python
"""Adds one workload's entry to profiles/r2_ncu_summary.json from an `ncu --set full` report, and writes the
per-launch table as text next to it.   usage: ncu_summary.py report.ncu-rep WORKLOAD "how it was captured" out.txt"""
import collections, csv, json, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, workload, how, out_txt = sys.argv[1:5]
# (a report too large to bring back from the GPU box is exported there: `ncu -i x.ncu-rep --page raw --csv > x.csv`)
raw = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def col(name, r):
    i = hdr.index(name)
    return float(r[i]) * SCALE.get(units[i], 1.0)


M = {"ms": "gpu__time_duration.sum", "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
     "lanes_per_inst": "smsp__thread_inst_executed_per_inst_executed.ratio", "warp_inst": "smsp__inst_executed.sum",
     "dram_read": "dram__bytes_read.sum", "dram_write": "dram__bytes_write.sum",
     "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1_hit_pct": "l1tex__t_sector_hit_rate.pct",
     "l2_hit_pct": "lts__t_sector_hit_rate.pct", "l1tex_pct": "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
     "fma_pipe_pct": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
     "alu_pipe_pct": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
     "xu_pipe_pct": "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
     "fp64_pipe_pct": "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
     "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
     "registers": "launch__registers_per_thread",
     "stall_long_scoreboard": "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"}
acc = collections.OrderedDict()
lines = [f"# {how}", f"# {os.path.basename(rep)}: one line per captured launch (times under ncu are serialised and cold)"]
ik = hdr.index("Kernel Name")
for r in data:
    name = re.sub(r"^void ", "", r[ik]).split("<")[0].split("(")[0].replace("rt::", "")
    v = {k: col(m, r) for k, m in M.items() if m in hdr}
    v["dram_bytes"] = v.pop("dram_read", 0.0) + v.pop("dram_write", 0.0)
    acc.setdefault(name, []).append(v)
    lines.append(f"{r[ik][:48]:48s} " + " ".join(f"{k}={v[k]:.4g}" for k in v))
entry = {"source": f"profiles/{os.path.basename(out_txt)} ({how})"}
for name, vs in acc.items():
    n = len(vs)
    e = {k: sum(x[k] for x in vs) / n for k in vs[0]}
    tot_ms = sum(x["ms"] for x in vs)
    entry[name] = {"launches_captured": n, "ms_per_launch": e["ms"], "issue_active_pct": e["issue_active_pct"],
                   # lanes per instruction weighted by the instructions of each launch
                   "lanes_per_inst": sum(x["lanes_per_inst"] * x["warp_inst"] for x in vs) / sum(x["warp_inst"] for x in vs),
                   "dram_bytes_per_launch": e["dram_bytes"], "dram_gbs": sum(x["dram_bytes"] for x in vs) / (tot_ms * 1e-3) / 1e9,
                   "dram_pct": e["dram_pct"], "l1_hit_pct": e["l1_hit_pct"], "l2_hit_pct": e["l2_hit_pct"],
                   "fma_pipe_pct": e.get("fma_pipe_pct"), "alu_pipe_pct": e.get("alu_pipe_pct"),
                   "xu_pipe_pct": e.get("xu_pipe_pct"), "fp64_pipe_pct": e.get("fp64_pipe_pct"),
                   "warps_active_pct": e.get("warps_active_pct"), "registers": e.get("registers"),
                   "stall_long_scoreboard_per_issue": e.get("stall_long_scoreboard")}
    lines.append(f"== {name}: " + json.dumps(entry[name]))
open(out_txt, "w").write("\n".join(lines) + "\n")
path = os.path.join(ROOT, "profiles", "r2_ncu_summary.json")
try:
    allw = json.load(open(path))
except (OSError, ValueError):
    allw = {}
allw[workload] = entry
json.dump(allw, open(path, "w"), indent=1)
print(json.dumps(entry, indent=1))
