import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as graft
rt = graft.load_package()
import oracle_binding as ob, scenes
tv, ti = scenes.teapot()
cam = rt.CameraVectors.default()
p = rt.FrameParams.make(160, 120, 0, 4)
inputs = rt.SceneInputs([(tv, ti)], [(0, rt.TYPE_MIRROR, rt.translation34(0, 0, 5))], None)
scene = rt.Scene(inputs, device_id=0)
prim = scene.primary_hits(cam, p)
ref = ob.OracleScene(inputs).render(cam, p, want=("primary",))["primary"]
i = (28 * 160 + 76) * 4
print("gpu", prim[i], "ref", ref[i])
