#!/usr/bin/env python3
"""usage: exp/launch_share.py launches.csv -> per-kernel share of a frame, averaged over the frames of the list.
The list is `ncu --metrics gpu__time_duration.sum --csv` of one bench.py command; a frame starts at k_raygen
(two k_raygen launches per frame when the frame is rendered as two sub-frames)."""
import csv, sys, collections, re
rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 14 and r[12] == "gpu__time_duration.sum"]
names = [re.sub(r"\(.*", "", r[4]) for r in rows]
ns = [float(r[14]) for r in rows]
first = names.index(next(n for n in names if "k_raygen" in n))
agg = collections.OrderedDict()
for n, t in zip(names[first:], ns[first:]):
    if not n.startswith(("rt::", "void rt::")): continue   # torch fill kernels of bench.py (L2 flush) are not the frame
    if ", 1>" in n or "k_fma_peak" in n: continue           # bench.py's one counting frame and its FP32 peak probe
    c, s = agg.get(n, (0, 0.0)); agg[n] = (c + 1, s + t)
ntrace = sum(c for n, (c, _) in agg.items() if "k_trace" in n)
frames = max(1, round(ntrace / 18))                         # 9 bounces x 2 sub-frames of closest-hit launches per frame
counting = round(sum(", 1>" in n and "k_trace" in n for n in names) / 18)
for n in list(agg):                                         # the untemplated kernels also ran in the counting frame(s)
    if "k_trace" not in n and "k_shadow" not in n:
        c, s = agg[n]; agg[n] = (c * frames / (frames + counting), s * frames / (frames + counting))
tot = sum(s for _, s in agg.values())
sub = round(sum("k_raygen" in n for n in names) / (frames + counting))   # one k_raygen per sub-frame
print(f"{frames} timed-path frames in the list ({sub} sub-frame(s) each); per frame:")
for n, (c, s) in agg.items():
    print(f"  {n:<28s} x{c / frames:5.1f} {s / frames / 1e6:8.3f} ms {100 * s / tot:5.1f} %")
print(f"  total {tot / frames / 1e6:.3f} ms (serialised, cold-cache launches)")
agg = collections.OrderedDict()
for n, t in zip(names[:first], ns[:first]):
    c, s = agg.get(n, (0, 0.0)); agg[n] = (c + 1, s + t)
print("before the first frame (build of both meshes, uploads):")
for n, (c, s) in agg.items():
    print(f"  {n[:60]:<60s} x{c:3d} {s / 1e3:9.1f} us")
