#!/usr/bin/env python3
"""usage: exp/make_traffic.py gpurun_out/r02_trav.ncu-rep  -> profiles/traffic.json + a per-launch table on stdout
Sums the ncu figures of the traversal launches of ONE C3 frame and tags them with the hash of the CUDA sources
(bench.py::source_hash), so that bench.py can tell whether they still describe the kernels it runs."""
import csv, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rep = sys.argv[1]
raw = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, data = rows[0], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
f = lambda r, h: float(r[ix[h]].replace(",", "")) if r[ix[h]] else 0.0
tot = {"dram": 0.0, "warp": 0.0, "thread": 0.0, "us": 0.0}
print("launch  kernel        us   lanes  issue%  warps%  L1hit%  L2hit%   rdMB   wrMB   Minst")
for n, r in enumerate(data):
    us = f(r, "gpu__time_duration.sum")
    unit = rows[1][ix["gpu__time_duration.sum"]]
    us = us * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "msecond": 1e3, "nsecond": 1e-3}.get(unit, 1.0)
    def b(h):
        v, u = f(r, h), rows[1][ix[h]]
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    rd, wr = b("dram__bytes_read.sum"), b("dram__bytes_write.sum")
    warp, thr = f(r, "smsp__inst_executed.sum"), f(r, "smsp__thread_inst_executed.sum") if "smsp__thread_inst_executed.sum" in ix else 0.0
    lanes = f(r, "smsp__thread_inst_executed_per_inst_executed.ratio")
    if not thr:
        thr = warp * lanes
    tot["dram"] += rd + wr; tot["warp"] += warp; tot["thread"] += thr; tot["us"] += us
    name = r[ix["Kernel Name"]].split("<")[0].split("::")[-1]
    print(f"{n:4d}  {name:10s} {us:8.1f} {lanes:6.1f} {f(r,'smsp__issue_active.avg.pct_of_peak_sustained_active'):7.1f} "
          f"{f(r,'sm__warps_active.avg.pct_of_peak_sustained_active'):7.1f} {f(r,'l1tex__t_sector_hit_rate.pct'):7.1f} {f(r,'lts__t_sector_hit_rate.pct'):7.1f} "
          f"{rd/1e6:7.1f} {wr/1e6:6.1f} {warp/1e6:7.1f}")
import bench
out = {"k_trace_dram_bytes_per_frame": int(tot["dram"]), "warp_instructions_per_frame": int(tot["warp"]),
       "thread_instructions_per_frame": int(tot["thread"]), "sm_count": 148, "sm_clock_hz": 1.965e9,
       "kernel_us_sum_under_ncu": round(tot["us"], 1), "launches": len(data), "source_hash": bench.source_hash(),
       "capture": os.path.basename(rep) + " (exp/ncu_traffic.sh: ncu --set full, the 36 k_trace + k_shadow launches of one C3 frame)"}
json.dump(out, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print(f"\nsum over the {len(data)} launches: {tot['us']/1e3:.3f} ms, DRAM {tot['dram']/1e9:.3f} GB, warp instructions {tot['warp']/1e9:.3f} G, "
      f"lanes {tot['thread']/tot['warp']:.2f}")
print("profiles/traffic.json written, source hash", out["source_hash"])
