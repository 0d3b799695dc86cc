# debug: fused path kernel vs oracle on the smoke scene: where do primary ids differ?
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as graft
rt = graft.load_package()
import oracle_binding as ob, scenes
inputs = scenes.smoke_scene(rt)
cam = rt.CameraVectors.default()
p = rt.FrameParams.make(160, 120, 6, 4)
scene = rt.Scene(inputs, device_id=0)
img = scene.render(cam, p)
st = scene.stats()
prim = scene.primary_hits(cam, p)
ref = ob.OracleScene(inputs).render(cam, p, want=("rgba8", "primary", "counts"))
rp = ref["primary"]
bad = (prim["instance"] != rp["instance"]) | (prim["primitive"] != rp["primitive"])
print("env", {k: v for k, v in os.environ.items() if k.startswith("RT_")})
print("stats", {k: st[k] for k in ("rays_primary", "rays_secondary", "rays_shadow", "n_miss", "n_hit", "stack_overflow", "max_tree_depth")}, "oracle", ref["counts"])
print("bad", bad.sum(), "of", bad.size, " img maxdiff", np.abs(img.astype(int) - ref["rgba8"].astype(int)).max())
idx = np.nonzero(bad)[0][:12]
for i in idx:
    pix, s = divmod(i, 4); y, x = divmod(pix, 160)
    print(f"  px ({x},{y}) s{s}: gpu inst {prim['instance'][i]:#x} prim {prim['primitive'][i]:#x} t {prim['t'][i]:.4f} | ref inst {rp['instance'][i]:#x} prim {rp['primitive'][i]:#x} t {rp['t'][i]:.4f}")
gi = prim["instance"][bad]; ri = rp["instance"][bad]
print("  gpu inst histogram of bad:", np.unique(gi, return_counts=True), " ref:", np.unique(ri, return_counts=True))
