# device vs host area-weighted normals: stand-in mesh (327 680 tri) and the flattened C5a grid (83.9 M tri)
import sys, os, time
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "tests"))
import numpy as np
import __graft_entry__ as graft
rt = graft.load_package()
v, i = rt.make_standin_armadillo(7)
for name, (vv, ii) in (("standin", (v, i)), ("c5a grid 16x16", rt.make_grid(v, i, 16, 16, 6.0))):
    rt.generate_normals(vv[:64], ii[:0])          # context warm-up
    t0 = time.perf_counter(); dev, ms = rt.generate_normals(vv, ii); t1 = time.perf_counter()
    host = rt.generate_normals_host(vv, ii); t2 = time.perf_counter()
    print(f"{name:16s} {ii.shape[0]/1e6:7.2f} M tri  device kernels {ms:8.2f} ms  (call with copies {1e3*(t1-t0):8.1f} ms)  "
          f"host loop {1e3*(t2-t1):9.1f} ms  equal {np.array_equal(dev, host)}", flush=True)
