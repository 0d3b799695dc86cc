#!/bin/sh
# usage: exp/run.sh v1 v2 ...  -> one summary line per variant in gpurun_out/exp.txt
cd "$(dirname "$0")/.."
for v in "$@"; do
  RT_B200_LIB=$PWD/exp/librt_$v.so python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/exp_$v.json 2> gpurun_out/exp_$v.err
  python - "$v" <<'PY' >> gpurun_out/exp.txt
import json,sys
v=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/exp_{v}.json"))
    r=d["roofline"]; s=d["stage_ms"]
    print(f"{v:12s} Mrays/s {d['value']:8.1f} frame {d['frame_ms']:7.3f} trace {s['trace']:7.3f} shadow {s['shadow']:6.3f} shade {s['shade_raygen_resolve']:6.3f} nodes {r['nodes_visited']/1e6:7.1f}M tris {r['tris_tested']/1e6:7.1f}M frac {r['frac']:.4f} build {d['build_ms']:.2f}")
except Exception as e:
    print(v, "FAILED", e, open(f"gpurun_out/exp_{v}.err").read()[-400:])
PY
done
cat gpurun_out/exp.txt
