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


def sea_sky_256():
    return np.load(os.path.join(GOLDEN, "sea_sky_256.npz"))["sky"]


def c1_scene(rt, sky=None):
    """BASELINE config C1: cube_scene.obj (corrected-normal policy) in the centre slot with the
    default CENTER_MESH_TYPE 1 (mirror, config.h:15), the orbiting slot filled with cube.obj at the default
    ORBITING_MESH_TYPE 0 (diffuse, config.h:16) at T(0,0,5) (main.cpp:1805-1808), skybox_texture_sea
    (config.h:4).  sky=None: the 256x256 fixture of that skybox (the 2048x2048 original is not a fixture)."""
    sv, si = golden_mesh("cube_scene")
    cv, ci = unit_cube()
    return rt.SceneInputs([(sv, si), (cv, ci)], [(0, rt.TYPE_MIRROR, rt.IDENTITY34),
                                                 (1, rt.TYPE_DIFFUSE, rt.translation34(0, 0, 5))],
                          sea_sky_256() if sky is None else sky)


def c1_bench_scene(rt, sky_size=None):
    """C1 for bench.py --workload c1 (the sea-sky fixture whatever sky size the other workloads use)."""
    return c1_scene(rt)


def c1_params(rt):
    """800x600 window (main.cpp:805), MAX_BOUNCE_COUNT 63, SAMPLES_PER_PIXEL 4 (config.h:26-27)."""
    return rt.FrameParams.make(800, 600, 63, 4)


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
        ("c1_cube_scene_800x600_d63_sky256", c1_scene(rt), c1_params(rt)),
    ]


def instanced_grid_scene(rt, nx=4, nz=3, level=2, sky_size=16):
    """Small stand-in for config C5b: one mesh, nx*nz instances on the XZ plane with alternating
    diffuse / refractive types plus a mirror teapot: enough instances for the top-level tree."""
    import math
    av, ai = rt.make_standin_armadillo(level)
    tv, ti = teapot()
    inst = []
    for iz in range(nz):
        for ix in range(nx):
            t = rt.translation34(6.0 * (ix - (nx - 1) / 2), -1.0 + 0.5 * ((ix + iz) % 3), -6.0 * iz)
            if (ix + iz) % 4 == 1:   # a few rotated instances
                a = 0.3 * (ix + 1)
                t = np.array([math.cos(a), 0, math.sin(a), t[3], 0, 1, 0, t[7], -math.sin(a), 0, math.cos(a), t[11]], np.float32)
            inst.append((0, rt.TYPE_DIFFUSE if (ix + iz) & 1 else rt.TYPE_REFRACTIVE, t))
    inst.append((1, rt.TYPE_MIRROR, rt.translation34(0, -4, 4)))
    return rt.SceneInputs([(av, ai), (tv, ti)], inst, rt.make_procedural_skybox(sky_size))


GRID_PITCH = 6.0   # 1.2 x the stand-in's width (radius 2.5)


def look_at_camera(rt, position, target):
    """Camera vectors for a camera at `position` looking at `target`, built the way the reference's
    Camera class derives them (camera.cpp:16-25): right = normalize(-f.z, 0, f.x), up = right x front."""
    p, t = np.asarray(position, np.float64), np.asarray(target, np.float64)
    f = (t - p) / np.linalg.norm(t - p)
    r = np.array([-f[2], 0.0, f[0]])
    r /= np.linalg.norm(r)
    u = np.cross(r, f)
    return rt.CameraVectors.make(p, r, u, f)


def c5_camera(rt, n=16):
    """Pulled back so the n x n grid fills the 2.5-focal-length view (half angle atan(1/2.5))."""
    half = 0.5 * n * GRID_PITCH
    return look_at_camera(rt, (0.0, 1.25 * half, 2.9 * half), (0.0, 0.0, 0.0))


def c5a_scene(rt, n=16, level=7, sky_size=2048):
    """BASELINE config C5a: n x n stand-in armadillos flattened into ONE mesh (16x16 at level 7 =
    83.9 M triangles), diffuse: stresses the LBVH build and node fetches from HBM."""
    av, ai = rt.make_standin_armadillo(level)
    gv, gi = rt.make_grid(av, ai, n, n, GRID_PITCH)
    return rt.SceneInputs([(gv, gi)], [(0, rt.TYPE_DIFFUSE, rt.IDENTITY34)], rt.make_procedural_skybox(sky_size))


def c5b_scene(rt, n=16, level=7, sky_size=2048):
    """BASELINE config C5b: one stand-in BLAS + n*n instance transforms, types alternating diffuse /
    refractive by (ix+iz)&1: exercises the top-level tree."""
    av, ai = rt.make_standin_armadillo(level)
    inst = []
    for iz in range(n):
        for ix in range(n):
            inst.append((0, rt.TYPE_REFRACTIVE if (ix + iz) & 1 else rt.TYPE_DIFFUSE,
                         rt.translation34((ix - 0.5 * (n - 1)) * GRID_PITCH, 0.0, (iz - 0.5 * (n - 1)) * GRID_PITCH)))
    return rt.SceneInputs([(av, ai)], inst, rt.make_procedural_skybox(sky_size))


def materials_for(rt, n):
    """n non-default materials: different exponents (one stays at the reference's 100,
    one far from it), coloured reflectances, different indices of refraction."""
    table = [rt.Material.make(ka=(0.3, 0.1, 0.1), kd=(0.9, 0.3, 0.2), ks=(0.5, 0.5, 0.9), ns=17.5, ni=1.33),
             rt.Material.make(ka=(0.05, 0.05, 0.2), kd=(0.1, 0.2, 0.9), ks=(1.0, 0.8, 0.2), ns=100.0, ni=2.4),
             rt.Material.make(ka=(0.2, 0.2, 0.0), kd=(0.8, 0.8, 0.1), ks=(0.3, 0.9, 0.3), ns=423.0, ni=1.05),
             rt.Material.make(ka=(0.0, 0.0, 0.0), kd=(1.0, 1.0, 1.0), ks=(0.0, 0.0, 0.0), ns=1.0, ni=1.52)]
    return [table[i % len(table)] for i in range(n)]
