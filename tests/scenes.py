"""Scene fixtures shared by tests, smoke() and bench.py.  Everything here is available on the
GPU box: meshes come from tests/golden/*.npz (packed by the reference's loader rules, see
tests/golden/make_golden.py) or are generated; nothing reads /root/reference."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_mesh(name):
    z = np.load(os.path.join(GOLDEN, f"{name}_packed.npz"))
    return z["vertices"].astype(np.float32), z["indices"].astype(np.uint32)


def unit_cube():
    return golden_mesh("cube")


def teapot():
    return golden_mesh("teapot")


def smoke_scene(rt):
    tv, ti = teapot()
    cv, ci = unit_cube()
    sky = rt.make_procedural_skybox(128)
    return rt.SceneInputs([(tv, ti), (cv, ci)],
                          [(0, rt.TYPE_MIRROR, rt.IDENTITY34),
                           (0, rt.TYPE_REFRACTIVE, rt.translation34(0, 0, 5)),
                           (1, rt.TYPE_DIFFUSE, rt.translation34(-4.5, 1.0, 2.0))], sky)


def c2_scene(rt, sky_size=2048):
    """BASELINE config C2: teapot mirror (centre) + teapot refractive at T(0,0,5)."""
    tv, ti = teapot()
    return rt.SceneInputs([(tv, ti)], [(0, rt.TYPE_MIRROR, rt.IDENTITY34),
                                       (0, rt.TYPE_REFRACTIVE, rt.translation34(0, 0, 5))],
                          rt.make_procedural_skybox(sky_size))


def c3_scene(rt, level=7, sky_size=2048):
    """BASELINE config C3: diffuse teapot in the centre slot, refractive armadillo (stand-in) in
    the orbiting slot at T(0,0,5) — the reference's own slot assignment (config.h:6-7) with the
    types the config names, so shadow rays exist (they fire only on diffuse hits)."""
    tv, ti = teapot()
    av, ai = rt.make_standin_armadillo(level)
    return rt.SceneInputs([(tv, ti), (av, ai)], [(0, rt.TYPE_DIFFUSE, rt.IDENTITY34),
                                                 (1, rt.TYPE_REFRACTIVE, rt.translation34(0, 0, 5))],
                          rt.make_procedural_skybox(sky_size))


def sea_sky_64():
    return np.load(os.path.join(GOLDEN, "sea_sky_64.npz"))["sky"]


def c1_like_scene(rt):
    """Small stand-in for config C1 (the reference's cube_scene.obj is malformed for its own
    loader): mirror teapot + diffuse cube, reference's sea skybox at 64x64."""
    tv, ti = teapot()
    cv, ci = unit_cube()
    return rt.SceneInputs([(tv, ti), (cv, ci)], [(0, rt.TYPE_MIRROR, rt.IDENTITY34),
                                                 (1, rt.TYPE_DIFFUSE, rt.translation34(0, 1, 5))], sea_sky_64())


def hash_cases(rt):
    """(name, inputs, params) of the oracle regression pins in tests/golden/oracle_hashes.json."""
    tv, ti = teapot()
    sky = sea_sky_64()
    return [
        ("c1_like_200x150_d63", c1_like_scene(rt), rt.FrameParams.make(200, 150, 63, 4)),
        ("c2_teapots_240x135_d4", rt.SceneInputs([(tv, ti)], [(0, rt.TYPE_MIRROR, rt.IDENTITY34),
                                                              (0, rt.TYPE_REFRACTIVE, rt.translation34(0, 0, 5))], sky),
         rt.FrameParams.make(240, 135, 4, 4)),
        ("smoke_160x120_d6", smoke_scene(rt), rt.FrameParams.make(160, 120, 6, 4)),
    ]
