#!/usr/bin/env python
"""Summarise `ncu --page source --csv --print-source sass` and `--page raw --csv` exports:
stall-reason totals, instruction mix, hottest SASS lines, headline metrics."""
import csv
import sys
from collections import Counter


def main(src_csv, raw_csv=None, top_n=25):
    rows = list(csv.reader(open(src_csv)))
    hdr, data = rows[1], rows[2:]
    isrc, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    cols = [c for c in hdr if c.startswith("stall_") and "Not Issued" not in c]
    tot = {c: sum(int(r[hdr.index(c)] or 0) for r in data) for c in cols}
    print(rows[0][1][:140])
    print("stalls:", {k: v for k, v in sorted(tot.items(), key=lambda kv: -kv[1]) if v})
    print("samples", sum(int(r[isamp] or 0) for r in data), "warp-inst", sum(int(r[iex] or 0) for r in data))
    mix = Counter()
    for r in data:
        t = r[isrc].split()
        op = t[0] if not t[0].startswith("@") else t[1]
        mix[op.split(".")[0]] += int(r[iex] or 0)
    print("mix:", mix.most_common(20))
    print("samples per 100 SASS lines:",
          [(s, sum(int(r[isamp] or 0) for r in data[s:s + 100])) for s in range(0, len(data), 100)])
    top = sorted(range(len(data)), key=lambda i: -int(data[i][isamp] or 0))[:top_n]
    for i in sorted(top):
        print(f"  {i:5d} samples={data[i][isamp]:>5s} exec={data[i][iex]:>8s}  {data[i][isrc][:90]}")
    if raw_csv:
        rr = list(csv.reader(open(raw_csv)))
        h, u, v = rr[0], rr[1], rr[2]
        for k in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
                  "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__grid_size",
                  "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
                  "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
                  "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
                  "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
                  "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sector_hit_rate.pct",
                  "dram__throughput.avg.pct_of_peak_sustained_elapsed"):
            if k in h:
                print(f"{k} [{u[h.index(k)]}] = {v[h.index(k)]}")


if __name__ == "__main__":
    main(*sys.argv[1:3])
