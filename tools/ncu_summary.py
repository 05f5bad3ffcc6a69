#!/usr/bin/env python
"""Summarise an .ncu-rep: per-launch key metrics (raw page) and, optionally, the hottest SASS lines.
usage: ncu_summary.py rep [--src KERNEL_INDEX] [--top N]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'lts__t_sector_hit_rate.pct', 'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active']
for r in rows[2:]:
    print(r[hdr.index('Kernel Name')][:80])
    for w in want:
        if w in hdr:
            print(f"   {w} = {r[hdr.index(w)]} {units[hdr.index(w)]}")
if "--src" in sys.argv:
    ki = int(sys.argv[sys.argv.index("--src") + 1])
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 30
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    secs = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
    si = secs[ki]
    h = rows[si + 1]
    body = rows[si + 2:secs[ki + 1]]
    isrc, ins = h.index("Source"), h.index("# Samples")
    stall = [i for i, x in enumerate(h) if x.startswith("stall_")]
    tot = sum(int(r[ins] or 0) for r in body if len(r) > ins)
    print(rows[si][1][:90], "total samples", tot)
    for k, r in sorted(enumerate(body), key=lambda kr: -int(kr[1][ins] or 0))[:top]:
        st = sorted(((h[i][6:], int(r[i] or 0)) for i in stall if (r[i] or '0') != '0' and 'Not' not in h[i]), key=lambda kv: -kv[1])[:2]
        print(f"{k:5d} {int(r[ins]):6d} {100 * int(r[ins]) / tot:5.1f}%  {r[isrc][:70]:70s} {st}")
