#!/usr/bin/env python3
"""usage: exp/launch_share.py launches.csv -> per-kernel share of a frame, averaged over the frames of the list.
The list is `ncu --metrics gpu__time_duration.sum --csv` of one bench.py command; a frame starts at k_raygen
(two k_raygen launches per frame when the frame is rendered as two sub-frames)."""
import csv, sys, collections, re
rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 14 and r[12] == "gpu__time_duration.sum"]
names = [re.sub(r"\(.*", "", r[4]) for r in rows]
ns = [float(r[14]) for r in rows]
first = names.index(next(n for n in names if "k_raygen" in n))
sub = 2 if "k_raygen" in names[first + 1] else 1   # two sub-frames: their k_raygen launches are adjacent
frames = sum("k_raygen" in n for n in names) // sub
agg = collections.OrderedDict()
for n, t in zip(names[first:], ns[first:]):
    if not n.startswith(("rt::", "void rt::")): continue   # torch fill kernels of bench.py (L2 flush) are not the frame
    c, s = agg.get(n, (0, 0.0)); agg[n] = (c + 1, s + t)
tot = sum(s for _, s in agg.values())
print(f"{frames} frames in the list ({sub} sub-frame(s) each); per frame:")
for n, (c, s) in agg.items():
    print(f"  {n:<28s} x{c // frames:3d} {s / frames / 1e6:8.3f} ms {100 * s / tot:5.1f} %")
print(f"  total {tot / frames / 1e6:.3f} ms (serialised, cold-cache launches)")
agg = collections.OrderedDict()
for n, t in zip(names[:first], ns[:first]):
    c, s = agg.get(n, (0, 0.0)); agg[n] = (c + 1, s + t)
print("before the first frame (build of both meshes, uploads):")
for n, (c, s) in agg.items():
    print(f"  {n[:60]:<60s} x{c:3d} {s / 1e3:9.1f} us")
