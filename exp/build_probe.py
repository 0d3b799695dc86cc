import sys, os
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "tests"))
import __graft_entry__ as graft
rt = graft.load_package()
import scenes
inputs = scenes.c5a_scene(rt, sky_size=64)
for k in range(3):
    sys.stderr.write(f"---- build {k}\n"); sys.stderr.flush()
    sc = rt.Scene(inputs, device_id=0)
    sys.stderr.write(f"build_ms {sc.stats()['build_ms']:.2f}\n")
    sc.close()
