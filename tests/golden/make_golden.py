"""Generates the committed fixtures under tests/golden/ from the reference checkout.
Run in the authoring container (needs /root/reference and oracle/_ref built by
oracle/ref_build.sh); the GPU box only reads the outputs.

  cube_packed.npz, teapot_packed.npz   vertex/index buffers exactly as the reference packs them
                                       (its own tiny_obj_loader.h + the rules of main.cpp:1638-1700)
  camera_golden.npz                    position/right/up/front and view matrices of the reference's
                                       unmodified Camera (src/camera.cpp) after a scripted input sequence
  sea_sky_64.npz                       the reference's skybox_texture_sea decoded by its stb_image.h,
                                       box-filtered to 64x64 per face (orientation / sampling tests)
  cube_scene_packed.npz                resources/cube_scene.obj under the corrected-normal policy (the
                                       reference's own loader reads past the end of `normals` on this file, so this
                                       one is packed by rth_load_obj, policy 1 = de-indexed per face corner)
  sea_sky_256.npz                      the same skybox box-filtered to 256x256 per face: config C1 on the GPU box
  oracle_hashes.json                   SHA-256 of oracle RGBA8 frames (regression pins of the oracle); the entry
                                       c1_cube_scene_800x600_d63_sky2048 is config C1 exactly as BASELINE.json states it
                                       (full 2048x2048 sea skybox, only renderable where the reference checkout exists)
"""
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pack(name):
    out = f"/tmp/{name}.bin"
    subprocess.check_call([os.path.join(ROOT, "oracle/_ref/ref_pack"), f"{REF}/resources/{name}.obj", out])
    raw = open(out, "rb").read()
    nv, nt = np.frombuffer(raw[:8], np.uint32)
    v = np.frombuffer(raw[8:8 + 24 * nv], np.float32).reshape(nv, 6)
    i = np.frombuffer(raw[8 + 24 * nv:], np.uint32).reshape(nt, 3)
    np.savez_compressed(os.path.join(HERE, f"{name}_packed.npz"), vertices=v, indices=i)
    return v, i


CAMERA_SCRIPT = [("mouse", 0.3, -0.2), ("move", 4, 2.5), ("move", 0, 1.25), ("mouse", -1.1, 0.9), ("move", 3, 0.75),
                 ("mouse", 0.0, 5.0), ("move", 5, 3.0), ("mouse", 2.0, -7.0), ("move", 1, 0.5), ("look", 2), ("move", 2, 1.0),
                 ("look", 5), ("mouse", 0.01, 0.02), ("move", 4, 10.0)]


def camera_golden():
    L = C.CDLL(os.path.join(ROOT, "oracle/_ref/libref_camera.so"))
    L.refcam_create.restype = C.c_void_p
    L.refcam_create.argtypes = [C.c_float] * 3
    L.refcam_move.argtypes = [C.c_void_p, C.c_int, C.c_float]
    L.refcam_mouse.argtypes = [C.c_void_p, C.c_float, C.c_float]
    L.refcam_look.argtypes = [C.c_void_p, C.c_int]
    L.refcam_get.argtypes = [C.c_void_p, C.c_void_p]
    L.refcam_view.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    cam = L.refcam_create(0.0, 0.0, 20.0)
    vecs, views = [], []

    def snap():
        o = np.zeros(12, np.float32)
        L.refcam_get(cam, o.ctypes.data)
        m = np.zeros(32, np.float32)
        L.refcam_view(cam, 0, m.ctypes.data)
        L.refcam_view(cam, 1, m[16:].ctypes.data)
        vecs.append(o.copy())
        views.append(m.copy())
    snap()
    for op in CAMERA_SCRIPT:
        if op[0] == "mouse":
            L.refcam_mouse(cam, op[1], op[2])
        elif op[0] == "move":
            L.refcam_move(cam, op[1], op[2])
        else:
            L.refcam_look(cam, op[1])
        snap()
    np.savez_compressed(os.path.join(HERE, "camera_golden.npz"), vectors=np.array(vecs), views=np.array(views))


def sea_sky():
    import __graft_entry__ as g
    rt = g.load_package()
    sky = rt.load_skybox_dir(f"{REF}/resources/skybox_texture_sea")  # reference's stb_image.h decode
    f = sky.shape[1] // 64
    small = sky.reshape(6, 64, f, 64, f, 4).astype(np.float64).mean((2, 4))
    np.savez_compressed(os.path.join(HERE, "sea_sky_64.npz"), sky=np.rint(small).astype(np.uint8))


def cube_scene():
    import __graft_entry__ as g
    rt = g.load_package()
    v, i, policy = rt.load_obj(f"{REF}/resources/cube_scene.obj")
    assert policy == 1
    np.savez_compressed(os.path.join(HERE, "cube_scene_packed.npz"), vertices=v, indices=i)
    return v, i


def sea_sky_256():
    import __graft_entry__ as g
    rt = g.load_package()
    sky = rt.load_skybox_dir(f"{REF}/resources/skybox_texture_sea")
    f = sky.shape[1] // 256
    small = sky.reshape(6, 256, f, 256, f, 4).astype(np.float64).mean((2, 4))
    np.savez_compressed(os.path.join(HERE, "sea_sky_256.npz"), sky=np.rint(small).astype(np.uint8))


def oracle_hashes():
    import __graft_entry__ as g
    rt = g.load_package()
    import oracle_binding as ob
    import scenes
    out = {}
    cam = rt.CameraVectors.default()
    for name, inputs, params in scenes.hash_cases(rt):
        img = ob.OracleScene(inputs).render(cam, params)["rgba8"]
        out[name] = hashlib.sha256(img.tobytes()).hexdigest()
    # config C1 with the reference's full-size skybox (not a fixture: 96 MiB); tests/test_c1.py re-renders it
    # wherever /root/reference exists
    full = scenes.c1_scene(rt, sky=rt.load_skybox_dir(f"{REF}/resources/skybox_texture_sea"))
    img = ob.OracleScene(full).render(cam, scenes.c1_params(rt))["rgba8"]
    out["c1_cube_scene_800x600_d63_sky2048"] = hashlib.sha256(img.tobytes()).hexdigest()
    json.dump(out, open(os.path.join(HERE, "oracle_hashes.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    for n in ("cube", "teapot"):
        v, i = pack(n)
        print(n, v.shape, i.shape)
    camera_golden()
    sea_sky()
    print("cube_scene", [a.shape for a in cube_scene()])
    sea_sky_256()
    oracle_hashes()
    print("golden fixtures written")
