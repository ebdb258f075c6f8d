#!/usr/bin/env python
"""Turns the ncu outputs in gpurun_out/ into the text summaries committed under profiles/."""
import collections, csv, json, re, subprocess, sys

def launches(csv_path, out_path, header):
    lines = [l for l in open(csv_path) if l.startswith('"')]
    rows = list(csv.reader(lines))
    hdr = rows[0]
    ik, iv = hdr.index('Kernel Name'), hdr.index('Metric Value')
    tot, n = collections.OrderedDict(), collections.Counter()
    for r in rows[1:]:
        k = re.sub(r'^void ', '', r[ik]).split('(')[0]
        tot[k] = tot.get(k, 0.0) + float(r[iv]); n[k] += 1
    T = sum(tot.values())
    with open(out_path, 'w') as f:
        f.write(header)
        f.write(f"{'kernel':58s} {'launches':>8s} {'total_ms':>10s} {'avg_us':>9s} {'share':>7s}\n")
        for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
            f.write(f"{k:58s} {n[k]:8d} {v / 1e6:10.3f} {v / n[k] / 1e3:9.1f} {v / T:7.3f}\n")
    return tot, n

KEEP = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_bytes.sum', 'lts__t_sector_hit_rate.pct', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_bytes.sum', 'l1tex__t_sector_hit_rate.pct', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio']

def full(rep_path, out_path, header, traffic_path=None):
    raw = subprocess.run(['ncu', '-i', rep_path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(out_path, 'w') as f:
        f.write(header)
        for k in KEEP:
            if k in hdr:
                i = hdr.index(k)
                f.write(f"{k} [{units[i]}]: " + " | ".join(r[i][:30] for r in data) + "\n")
    if traffic_path:
        ik, ir, iw, it = (hdr.index(x) for x in ('Kernel Name', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__time_duration.sum'))
        scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
        acc = collections.defaultdict(list)
        for r in data:
            name = re.sub(r'^void ', '', r[ik]).split('<')[0].split('(')[0]
            acc[name].append((float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]))
        json.dump({k: {"dram_bytes_per_launch": sum(v) / len(v), "launches_captured": len(v)} for k, v in acc.items()},
                  open(traffic_path, 'w'), indent=1)

if __name__ == '__main__':
    tag = sys.argv[1] if len(sys.argv) > 1 else 'r1'
    launches('gpurun_out/r1_bench_launches.csv', f'profiles/{tag}_bench_launches.txt',
             "# ncu --metrics gpu__time_duration.sum --clock-control none -c 400 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu\n"
             "# (stress_1M_4k on one B200; per-launch times under ncu are serialised and cold: read the SHARES)\n")
    full('gpurun_out/r1_bvh_full_d.ncu-rep', f'profiles/{tag}_bvh_kernels_ncu_full_d.txt',
         "# ncu --set full --clock-control none --import-source on -k regex:bvh_(trace|shadow|shade)_kernel -s 40 -c 8\n"
         "#   python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu      (stress_1M_4k, one B200)\n"
         "# State: closest hit on the 4-wide quantised LBVH, any hit on the 8-wide one (256-bit node loads), persistent\n"
         "#        while-while walks, shared-memory stack, shadow chains, two streams; occluder hints and the fused trace\n"
         "#        kernel off (DESIGN.md 5.5).  Full-frame launches: the builds without work sharing.\n",
         f'profiles/{tag}_traffic.json')
