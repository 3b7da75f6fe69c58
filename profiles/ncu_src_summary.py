#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump (first captured launch, SASS view): executed
warp-instructions per opcode, stall samples per SASS line, and a compact listing.
usage: ncu_src_summary.py <src.csv> [trials] [--listing out.txt]"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
trials = float(sys.argv[2]) if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else 65536.0
listing = sys.argv[sys.argv.index("--listing") + 1] if "--listing" in sys.argv else None
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
blk = rows[starts[0]:starts[1]] if len(starts) > 1 else rows[starts[0]:]
print("kernel:", blk[0][1][:100])
hdr = blk[1]
ix = {h: i for i, h in enumerate(hdr)}
STALLS = ["stall_long_sb", "stall_math", "stall_wait", "stall_short_sb", "stall_not_selected", "stall_dispatch",
          "stall_branch_resolving", "stall_mio", "stall_barrier", "stall_lg", "stall_selected"]
ops = collections.Counter(); samples = collections.Counter(); excess = collections.Counter()
stall_tot = collections.Counter()
lines = []
for r in blk[2:]:
    if len(r) < len(hdr):
        continue
    src = r[ix["Source"]]
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
    op = m.group(2) if m else "?"
    base = op.split(".")[0]
    n = float(r[ix["Instructions Executed"]] or 0)
    s = float(r[ix["# Samples"]] or 0)
    ex = float(r[ix["L1 Wavefronts Shared Excessive"]] or 0)
    wf = float(r[ix["L1 Wavefronts Shared"]] or 0)
    ops[base] += n; samples[base] += s; excess[base] += ex
    st = {k: float(r[ix[k]] or 0) for k in STALLS if k in ix}
    for k, v in st.items():
        stall_tot[k] += v
    lines.append((n, s, ex, wf, r[ix["Address"]], src, st))
tot = sum(ops.values()); tots = sum(samples.values())
print(f"total warp-instructions {tot:.0f} = {tot/trials:.1f} per trial; stall samples {tots:.0f}")
print("stall mix: " + "  ".join(f"{k[6:]} {100*v/max(tots,1):.1f}%" for k, v in stall_tot.most_common()))
print(f"{'opcode':12s} {'inst/trial':>10s} {'%inst':>6s} {'%samples':>8s} {'excess wf/trial':>16s}")
for op, n in ops.most_common(24):
    print(f"{op:12s} {n/trials:10.1f} {100*n/tot:6.1f} {100*samples[op]/max(tots,1):8.1f} {excess[op]/trials:16.1f}")
print("\nhottest SASS lines by stall samples:")
for n, s, ex, wf, a, src, st in sorted(lines, key=lambda t: -t[1])[:22]:
    top = max(st, key=st.get)[6:] if st else ""
    print(f"{100*s/max(tots,1):5.1f}%  inst/trial {n/trials:7.1f}  {top:12s} {src[:80]}")
if listing:
    with open(listing, "w") as f:
        for i, (n, s, ex, wf, a, src, st) in enumerate(lines):
            top = max(st, key=st.get)[6:] if st and s else ""
            f.write(f"{i:4d} {n/trials:8.2f} {100*s/max(tots,1):5.2f}% wf {wf/trials:7.2f} ex {ex/trials:6.2f} {top:10s} {src[:100]}\n")
