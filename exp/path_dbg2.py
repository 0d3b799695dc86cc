import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as graft
rt = graft.load_package()
import oracle_binding as ob, scenes
tv, ti = scenes.teapot()
sky = rt.make_procedural_skybox(64)
cam = rt.CameraVectors.default()
p = rt.FrameParams.make(160, 120, 6, 4)
cases = {
  "one@origin mirror": [(0, rt.TYPE_MIRROR, rt.IDENTITY34)],
  "one@z5 mirror": [(0, rt.TYPE_MIRROR, rt.translation34(0, 0, 5))],
  "one@z5 refr": [(0, rt.TYPE_REFRACTIVE, rt.translation34(0, 0, 5))],
  "one@z5 diffuse": [(0, rt.TYPE_DIFFUSE, rt.translation34(0, 0, 5))],
  "two: origin,z5": [(0, rt.TYPE_MIRROR, rt.IDENTITY34), (0, rt.TYPE_MIRROR, rt.translation34(0, 0, 5))],
  "two: z5,origin": [(0, rt.TYPE_MIRROR, rt.translation34(0, 0, 5)), (0, rt.TYPE_MIRROR, rt.IDENTITY34)],
}
for name, inst in cases.items():
    inputs = rt.SceneInputs([(tv, ti)], inst, sky)
    scene = rt.Scene(inputs, device_id=0)
    img = scene.render(cam, p); st = scene.stats()
    prim = scene.primary_hits(cam, p)
    ref = ob.OracleScene(inputs).render(cam, p, want=("rgba8", "primary", "counts"))
    rp = ref["primary"]
    bad = (prim["instance"] != rp["instance"]) | (prim["primitive"] != rp["primitive"])
    print(f"{name:20s} bad {bad.sum():6d} imgdiff {np.abs(img.astype(int) - ref['rgba8'].astype(int)).max():3d} sec {st['rays_secondary']} / {ref['counts']['secondary']} shadow {st['rays_shadow']} / {ref['counts']['shadow']}  ref inst of bad {np.unique(rp['instance'][bad], return_counts=True)}")
    scene.close()
