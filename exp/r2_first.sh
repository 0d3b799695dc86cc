#!/bin/sh
# round 2, first GPU contact of the fused path kernel: parity tests, then C3 with and without it
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/r2_tests.log
for f in 1 0; do
  RT_FUSED=$f timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_c3_fused$f.json 2> gpurun_out/r2_c3_fused$f.err; echo "bench fused=$f rc=$?"
  python - gpurun_out/r2_c3_fused$f.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); s=d["stage_ms"]; r=d["roofline"]
    print(f"  Mrays/s {d['value']:8.1f} frame {d['frame_ms']:7.3f} trace {s['trace']:7.3f} shadow {s['shadow']:6.3f} shade {s['shade_raygen_resolve']:6.3f} nodes {r['nodes_visited']/1e6:7.1f}M tris {r['tris_tested']/1e6:7.1f}M launches {d['gpu_launches']} e2e {d['e2e']['ms_per_step']:.3f}")
except Exception as e:
    print("  FAILED", e)
PY
done
