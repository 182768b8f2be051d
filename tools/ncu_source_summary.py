#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --csv --print-source=sass` of one kernel: stall samples and executed
instructions per 200-instruction region, and the single instructions with the most stall samples.
Usage: python tools/ncu_source_summary.py source.csv [label]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
label = sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]
hdr, data = rows[1], rows[2:]
isrc, iss, iex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
tot_s, tot_e = sum(int(r[iss]) for r in data), sum(int(r[iex]) for r in data)
print(f"== {label}: {rows[0][1]}")
print(f"stall samples {tot_s}, warp instructions executed {tot_e}, SASS instructions {len(data)}")
print("region (SASS index) | stall samples % | executed instructions % | first instruction")
for k in range(0, len(data), 200):
    ch = data[k:k + 200]
    s, e = sum(int(r[iss]) for r in ch), sum(int(r[iex]) for r in ch)
    if s / tot_s > 0.01 or e / tot_e > 0.01:
        print(f"{k:6d} | {100 * s / tot_s:5.1f} | {100 * e / tot_e:5.1f} | {ch[0][isrc].strip()[:48]}")
print("top instructions by stall samples: index | instruction | samples (% of all) | executions | main reason")
reasons = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
for r in sorted(data, key=lambda r: -int(r[iss]))[:12]:
    why = max(reasons, key=lambda i: int(r[i] or 0))
    print(f"{data.index(r):6d} | {r[isrc].strip()[:44]:44s} | {r[iss]} ({100 * int(r[iss]) / tot_s:.1f}) | {r[iex]} | {hdr[why]}")
