#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: hottest SASS instructions with their dominant stall reason."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
items, tot = [], 0
for r in rows[2:]:
    try:
        s = float(r[ix["Warp Stall Sampling (All Samples)"]])
    except Exception:
        continue
    tot += s
    items.append((s, r))
print("total samples", int(tot))
agg = {h: 0.0 for h in stalls}
for s, r in items:
    for h in stalls:
        try: agg[h] += float(r[ix[h]])
        except Exception: pass
print("stall mix:", {k: round(v / max(tot, 1), 3) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
items.sort(key=lambda x: -x[0])
for s, r in items[:top]:
    dom = max(stalls, key=lambda h: float(r[ix[h]] or 0))
    print(f"{int(s):7d} {100*s/tot:5.1f}%  exec={r[ix['Instructions Executed']]:>10}  {dom:22s} | {r[ix['Source']].strip()[:100]}")
