#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: executed warp-instructions per opcode and per
kernel section, stall samples per SASS line.  usage: ncu_src_summary.py <src.csv> [trials]"""
import csv, sys, re, collections
rows = list(csv.reader(open(sys.argv[1])))
trials = float(sys.argv[2]) if len(sys.argv) > 2 else 65536.0
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
ops = collections.Counter(); samples = collections.Counter(); excess = collections.Counter()
lines = []
for r in rows[2:]:
    if len(r) < len(hdr): continue
    src = r[ix["Source"]]
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
    op = m.group(2) if m else "?"
    base = op.split(".")[0]
    n = float(r[ix["Instructions Executed"]] or 0)
    s = float(r[ix["# Samples"]] or 0)
    ex = float(r[ix["L1 Wavefronts Shared Excessive"]] or 0)
    ops[base] += n; samples[base] += s; excess[base] += ex
    lines.append((n, s, ex, r[ix["Address"]], src))
tot = sum(ops.values()); tots = sum(samples.values())
print(f"total warp-instructions {tot:.0f} = {tot/trials:.1f} per trial; stall samples {tots:.0f}")
print(f"{'opcode':12s} {'inst/trial':>10s} {'%inst':>6s} {'%samples':>8s} {'excess wavefronts/trial':>24s}")
for op, n in ops.most_common(28):
    print(f"{op:12s} {n/trials:10.1f} {100*n/tot:6.1f} {100*samples[op]/max(tots,1):8.1f} {excess[op]/trials:24.1f}")
print("\nhottest SASS lines by stall samples:")
for n, s, ex, a, src in sorted(lines, key=lambda t: -t[1])[:25]:
    print(f"{100*s/max(tots,1):5.1f}%  inst/trial {n/trials:7.1f}  excess {ex/trials:6.1f}  {src[:90]}")
