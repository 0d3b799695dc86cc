# SURVEY 8(f) rank 3, BVH cache file: would loading a saved acceleration structure beat rebuilding it?
# C5a (83.9 M triangles): rt_scene_build from the caller's host buffers (upload + build) against the upload of an
# image of the finished nodes + leaf triangles of the same scene (what a cache hit has to do after reading the file).
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import __graft_entry__ as g
rt = g.load_package()
import scenes
for name, inputs in (("C3 (330 k triangles)", scenes.c3_scene(rt, sky_size=64)), ("C5a (83.9 M triangles)", scenes.c5a_scene(rt, sky_size=64))):
    in_bytes = sum(v.nbytes + i.nbytes for v, i in inputs.meshes)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(2):
        t0 = time.perf_counter()
        sc = rt.Scene(inputs, device_id=0)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        st = sc.stats()
        sc.close()
        best = min(best, dt)
    bvh_bytes = st["bvh_bytes"]
    img = np.empty(bvh_bytes, np.uint8)          # the cache image as a file read would leave it: pageable host memory
    img[::4096] = 1
    dev = torch.empty(bvh_bytes, dtype=torch.uint8, device="cuda")
    ups = []
    for rep in range(2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        dev.copy_(torch.from_numpy(img))
        torch.cuda.synchronize()
        ups.append(time.perf_counter() - t0)
    pinned = torch.from_numpy(img).pin_memory()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    dev.copy_(pinned)
    torch.cuda.synchronize()
    up_pin = time.perf_counter() - t0
    print(f"{name}: inputs {in_bytes/1e9:.3f} GB -> rt_scene_build (upload + build, host wall) {best*1e3:.1f} ms, of which build kernels {st['build_ms']:.1f} ms; "
          f"BVH image {bvh_bytes/1e9:.3f} GB -> upload from pageable memory {min(ups)*1e3:.1f} ms, from pinned memory {up_pin*1e3:.1f} ms "
          f"(+ vertices/indices {in_bytes/1e9:.3f} GB still needed for shading, + the file read)", flush=True)
    del dev, pinned, img
