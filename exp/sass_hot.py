#!/usr/bin/env python3
"""usage: ncu -i X.ncu-rep --page source --csv --print-source sass > src.csv; exp/sass_hot.py src.csv [kernel index]
Per kernel of the report: stall mix, lane-occupancy histogram of the executed warp instructions, hottest SASS lines."""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1], errors="replace")))
kern = []
for i, r in enumerate(rows):
    if r and r[0] == "Kernel Name": kern.append(i)
kern.append(len(rows))
want = int(sys.argv[2]) if len(sys.argv) > 2 else None
for k in range(len(kern) - 1):
    if want is not None and k != want: continue
    name = rows[kern[k]][1]; hdr = rows[kern[k] + 1]; data = [r for r in rows[kern[k] + 2:kern[k + 1]] if len(r) == len(hdr)]
    ix = {h: i for i, h in enumerate(hdr)}
    f = lambda r, h: float(r[ix[h]] or 0)
    ti = sum(f(r, "Instructions Executed") for r in data); ts = sum(f(r, "# Samples") for r in data)
    print(f"== kernel {k}: {name[:30]}  warp insts {ti/1e6:.1f}M  samples {ts:.0f}  sass lines {len(data)}")
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = {h: sum(f(r, h) for r in data) for h in stalls}
    print("  stalls: " + "  ".join(f"{h[6:]} {100*v/ts:.1f}%" for h, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
    b = collections.Counter()
    for r in data:
        n = f(r, "Instructions Executed")
        if n: b[int(f(r, "Avg. Threads Executed") // 4) * 4] += n
    print("  lanes:  " + "  ".join(f"{k2}-{k2+3}: {100*b[k2]/ti:.1f}%" for k2 in sorted(b)))
    if want is not None:
        for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:45]:
            print(f'  {r[ix["# Samples"]]:>6s} {f(r,"Instructions Executed")/1e3:8.0f}k lanes {f(r,"Avg. Threads Executed"):5.1f}  {r[1].strip()[:64]:64s} lsb {r[ix["stall_long_sb"]]} ssb {r[ix["stall_short_sb"]]} wait {r[ix["stall_wait"]]} br {r[ix["stall_branch_resolving"]]} nosel {r[ix["stall_not_selected"]]}')
