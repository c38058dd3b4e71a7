#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --csv` (SASS view): samples and executed instructions by
opcode, stall reasons overall, and the hottest instructions.  Usage: ncu_sass_summary.py file.csv [ntop]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
def f(r, h):
    try: return float(r[ix[h]])
    except Exception: return 0.0
tot_s = sum(f(r, "# Samples") for r in data); tot_i = sum(f(r, "Instructions Executed") for r in data)
print(f"kernel: {rows[0][1][:70]}  SASS lines {len(data)}  samples {tot_s:.0f}  warp-instr {tot_i:.0f}")
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("stall reasons (all samples):", ", ".join(f"{h[6:]} {100*sum(f(r,h) for r in data)/max(tot_s,1):.1f}%" for h in sorted(stalls, key=lambda h: -sum(f(r,h) for r in data))[:8]))
byop_s, byop_i = collections.Counter(), collections.Counter()
for r in data:
    op = r[ix["Source"]].split()
    op = [t for t in op if not t.startswith("@")]
    name = op[0].split(".")[0] if op else "?"
    if name in ("F2F", "MUFU", "I2F", "F2I"): name = op[0]
    byop_s[name] += f(r, "# Samples"); byop_i[name] += f(r, "Instructions Executed")
print("by opcode: op  %instr  %samples")
for op, n in byop_i.most_common(18):
    print(f"  {op:14s} {100*n/tot_i:6.2f} {100*byop_s[op]/max(tot_s,1):6.2f}")
print("hottest instructions:")
for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:ntop]:
    top = max(stalls, key=lambda h: f(r, h))
    print(f"  {r[ix['Address']][-5:]} {100*f(r,'# Samples')/tot_s:5.2f}% {top[6:]:10s} {r[ix['Source']][:80]}")
