#!/bin/sh
# usage: exp/sweep_path.sh "ENV=.. ENV=.." ...   one C3 bench line per setting
cd "$(dirname "$0")/.."
for cfg in "$@"; do
  env $cfg timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/sw.json 2> gpurun_out/sw.err
  python - "$cfg" <<'PY'
import json,sys
try:
    d=json.load(open("gpurun_out/sw.json")); s=d["stage_ms"]; r=d["roofline"]
    print(f"{sys.argv[1]:50s} frame {d['frame_ms']:7.3f} trace {s['trace']:7.3f} shade {s['shade_raygen_resolve']:6.3f} nodes {r['nodes_visited']/1e6:7.1f}M tris {r['tris_tested']/1e6:6.1f}M e2e {d['e2e']['ms_per_step']:.3f}", flush=True)
except Exception as e:
    print(sys.argv[1], "FAILED", e, open("gpurun_out/sw.err").read()[-300:])
PY
done
