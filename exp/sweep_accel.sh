#!/bin/sh
# usage: exp/sweep_accel.sh workload accel...   one bench line per accel flavour
cd "$(dirname "$0")/.."
W=$1; shift
for a in "$@"; do
  timeout 900 python bench.py --workload $W --accel $a --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/acc_${W}_$a.json 2> gpurun_out/acc_${W}_$a.err
  python - $W $a <<'PY'
import json,sys
w,a=sys.argv[1:3]
try:
    d=json.loads([l for l in open(f"gpurun_out/acc_{w}_{a}.json") if l.startswith("{")][-1]); s=d["stage_ms"]; r=d["roofline"]
    print(f"{w} accel {a}: frame {d['frame_ms']:7.3f} trace {s['trace']:7.3f} shadow {s['shadow']:6.3f} nodes {r['nodes_visited']/1e6:7.1f}M tris {r['tris_tested']/1e6:6.1f}M build {d['build_ms']:.1f} ms bvh {d['bvh']['bytes']/1e6:.0f} MB", flush=True)
except Exception as e:
    print(w, a, "FAILED", e, open(f"gpurun_out/acc_{w}_{a}.err").read()[-400:])
PY
done
