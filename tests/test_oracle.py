"""Pins of the CPU oracle itself.  The reference ships no tests or golden vectors for this path
(parity unpinned), so the oracle is anchored on analytic known answers, on a double-precision
re-check of its FP32 intersector, on brute-force vs BVH identity and on committed hashes."""
import hashlib
import json
import math
import os

import numpy as np
import pytest

import scenes


def solid_sky(colors, size=8):
    sky = np.zeros((6, size, size, 4), np.uint8)
    for f, c in enumerate(colors):
        sky[f, :, :, :3] = c
        sky[f, :, :, 3] = 255
    return sky


FACE_COLORS = [(255, 0, 0), (0, 255, 0), (0, 0, 255), (255, 255, 0), (255, 0, 255), (0, 255, 255)]


def test_det_sin_accuracy(ob):
    rng = np.random.default_rng(1)
    xs = np.concatenate([rng.uniform(-4, 4, 2000), rng.uniform(-3e5, 3e5, 4000)]).astype(np.float32)
    got = np.array([ob.det_sin(x) for x in xs], np.float64)
    want = np.sin(xs.astype(np.float64))
    assert np.max(np.abs(got - want)) < 6.1e-8  # within one float ulp of the exact value near |sin| = 1
    assert ob.det_sin(0.0) == 0.0


def test_random_hash_matches_formula(ob):
    # shader.rgen:57-59 evaluated in float64 from the same float32 argument
    for (x, y, seed) in [(0, 0, 4), (17, 3, 4.5), (799, 599, 7), (3839, 2159, 7.5)]:
        d = np.float32(np.float32(x) * np.float32(12.9898)) if False else None
        got = ob.random_hash(x, y, seed)
        assert 0.0 <= got < 1.0
    # exact re-computation with numpy float32 + math.fma-free path for one case
    x, y, seed = np.float32(12), np.float32(34), np.float32(5)
    prod = np.float32(x * np.float32(12.9898))
    d = np.float32(float(np.float64(y) * np.float64(np.float32(78.233)) + np.float64(prod)))  # fma = one rounding
    a = np.float32(float(np.float64(np.float32(1113.1)) * np.float64(seed) + np.float64(d)))
    r = np.float32(np.float32(ob.det_sin(a)) * np.float32(43758.5453))
    assert ob.random_hash(12, 34, 5) == np.float32(r - np.floor(r))


def test_cube_known_answer(rt, ob):
    """cube.obj [-1,1]^3 at the origin, default camera (0,0,20) looking down -Z: rays inside the
    silhouette hit the z=+1 face, primitives 0 (x>y) and 1 (y>x), at t = 19/|d.z|."""
    cv, ci = scenes.unit_cube()
    sc = ob.OracleScene(rt.SceneInputs([(cv, ci)], [(0, rt.TYPE_DIFFUSE, rt.IDENTITY34)], None))
    rng = np.random.default_rng(2)
    pts = rng.uniform(-0.95, 0.95, (500, 2))
    pts = pts[np.abs(pts[:, 0] - pts[:, 1]) > 1e-3]
    o = np.tile(np.array([0, 0, 20], np.float32), (len(pts), 1))
    tgt = np.concatenate([pts, np.ones((len(pts), 1))], 1).astype(np.float32)
    d = tgt - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    for brute in (False, True):
        h = sc.trace_rays(o, d, brute=brute)
        assert np.all(h["instance"] == 0)
        assert np.array_equal(h["primitive"], np.where(pts[:, 0] > pts[:, 1], 0, 1).astype(np.uint32))
        assert np.allclose(h["t"], 19.0 / np.abs(d[:, 2]), rtol=2e-6)
    # rays outside the silhouette miss
    out = np.array([[1.5, 0.2, 1.0], [-1.2, -1.3, 1.0]], np.float32) - o[:2]
    out /= np.linalg.norm(out, axis=1, keepdims=True)
    assert np.all(sc.trace_rays(o[:2], out)["instance"] == rt.MISS)


def test_fp32_intersector_against_fp64(rt, ob):
    tv, ti = scenes.teapot()
    sc = ob.OracleScene(rt.SceneInputs([(tv, ti)], [(0, 1, rt.IDENTITY34), (0, 2, rt.translation34(0, 0, 5))], None))
    rng = np.random.default_rng(3)
    n = 400
    o = np.tile(np.array([0, 0, 20], np.float32), (n, 1))
    tgt = rng.uniform([-3, 0, -2], [3.4, 3.1, 7], (n, 3)).astype(np.float32)
    d = tgt - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    h = sc.trace_rays(o, d)
    agree = 0
    for k in range(n):
        i64, p64, t64 = sc.trace_f64(o[k], d[k])
        if h["instance"][k] == i64 and h["primitive"][k] == p64:
            agree += 1
            if i64 != rt.MISS:
                assert abs(h["t"][k] - t64) <= 4e-6 * t64
    assert agree >= n - 2  # residual only at exact edge hits


def test_bvh_equals_brute_force(rt, ob):
    inputs = scenes.smoke_scene(rt)
    sc = ob.OracleScene(inputs)
    p = rt.FrameParams.make(96, 72, 6, 2)
    cam = rt.CameraVectors.default()
    a = sc.render(cam, p, brute=False, want=("rgba32f", "primary", "counts"))
    b = sc.render(cam, p, brute=True, want=("rgba32f", "primary", "counts"))
    assert np.array_equal(a["rgba32f"], b["rgba32f"])
    for f in ("instance", "primitive", "t", "u", "v"):
        assert np.array_equal(a["primary"][f], b["primary"][f])
    for k in ("primary", "secondary", "shadow", "miss", "hit"):
        assert a["counts"][k] == b["counts"][k]
    c = a["counts"]
    assert c["primary"] == 96 * 72 * 2 and c["miss"] + c["hit"] == c["primary"] + c["secondary"]


def test_skybox_orientation(rt, ob):
    """Layer order +X,-X,+Y,-Y,+Z,-Z and the lookup vector (d.x, d.y, -d.z) of shader.rgen:92:
    the default camera looking down -Z must see layer 4 (front.jpg)."""
    cv, ci = scenes.unit_cube()
    inputs = rt.SceneInputs([(cv, ci)], [(0, 0, rt.translation34(0, 0, 1000))], solid_sky(FACE_COLORS))
    sc = ob.OracleScene(inputs)
    p = rt.FrameParams.make(9, 9, 1, 1)
    expect = {rt.FORWARD: 4, rt.BACKWARD: 5, rt.RIGHT: 0, rt.LEFT: 1, rt.UP: 2, rt.DOWN: 3}
    for look, layer in expect.items():
        cam = rt.Camera((0, 0, 0))
        cam.look(look)
        img = sc.render(cam.vectors(), p)["rgba8"]
        assert tuple(img[4, 4, :3]) == FACE_COLORS[layer], (look, layer)
    # texel addressing inside a face: +Z face, s grows with +x, t grows with -y
    sky = solid_sky([(0, 0, 0)] * 6, 4)
    sky[4, 0, 3, :3] = (255, 255, 255)  # row 0 (top), column 3 (right)
    sc2 = ob.OracleScene(rt.SceneInputs([(cv, ci)], [(0, 0, rt.translation34(0, 0, 1000))], sky))
    v = sc2.sky_sample(np.array([0.75, 0.75, 1.0], np.float32))
    assert np.allclose(v, 1.0) and np.allclose(sc2.sky_sample(np.array([-0.75, 0.75, 1.0], np.float32)), 0.0)


def test_slab_refraction_and_mirror_known_answers(rt, ob):
    """A ray through two parallel faces of a refractive cube leaves with its entry direction, so
    its colour equals the skybox in the primary direction; a mirror face z=1 reflects d to
    (dx,dy,-dz).  Both follow from shader.rgen:132-165 and pin the scatter code analytically."""
    cv, ci = scenes.unit_cube()
    sky = scenes.sea_sky_64()
    p = rt.FrameParams.make(64, 64, 4, 1)
    cam = rt.CameraVectors.default()
    empty = ob.OracleScene(rt.SceneInputs([(cv, ci)], [(0, 0, rt.translation34(0, 0, 1000))], sky))
    glass = ob.OracleScene(rt.SceneInputs([(cv, ci)], [(0, rt.TYPE_REFRACTIVE, rt.IDENTITY34)], sky))
    a = empty.render(cam, p, want=("rgba32f",))["rgba32f"]
    b = glass.render(cam, p, want=("rgba32f", "primary"))
    hit = (b["primary"]["instance"] == 0).reshape(64, 64)
    inner = np.zeros_like(hit)
    inner[30:34, 30:34] = True     # rays near the axis cross the front and back faces
    assert hit[inner].all()
    assert np.allclose(a[inner], b["rgba32f"][inner], atol=2e-3)
    mirror = ob.OracleScene(rt.SceneInputs([(cv, ci)], [(0, rt.TYPE_MIRROR, rt.IDENTITY34)], sky))
    m = mirror.render(cam, p, want=("rgba32f",))["rgba32f"]
    # centre pixel: d ~ (0,0,-1) reflects to (0,0,+1); lookup vector (0,0,-1) = layer 5 (back)
    want = mirror.sky_sample(np.array([0.0, 0.0, -1.0], np.float32))
    assert np.allclose(m[32, 32, :3], want, atol=0.03)


def test_phong_known_answer(rt, ob):
    """Front-facing diffuse triangle, unoccluded light: ambient + pow(0.9,i)*(kd*N.L + ks*(N.H)^100)
    (shader.rgen:116-128), checked in float64 for sample 0."""
    v = np.array([[-50, -50, 0, 0, 0, 1], [50, -50, 0, 0, 0, 1], [0, 50, 0, 0, 0, 1]], np.float32)
    inputs = rt.SceneInputs([(v, np.array([[0, 1, 2]], np.uint32))], [(0, 0, rt.IDENTITY34)], None)
    sc = ob.OracleScene(inputs)
    p = rt.FrameParams.make(1, 1, 0, 1)
    img = sc.render(rt.CameraVectors.default(), p, want=("rgba32f", "primary"))
    d = np.array([0, 0, 20.0])
    h = img["primary"][0]
    assert h["instance"] == 0
    # reconstruct the hit point from barycentrics (shader.rchit) and evaluate Phong
    P = v[0, :3] * (1 - h["u"] - h["v"]) + v[1, :3] * h["u"] + v[2, :3] * h["v"]
    dirv = (P - d) / np.linalg.norm(P - d)
    L = np.array([5, 5, 5.0]) - P
    L /= np.linalg.norm(L)
    Hh = (L - dirv) / np.linalg.norm(L - dirv)
    N = np.array([0, 0, 1.0])
    col = np.array([0.08, 0.24, 0.08]) + np.array([0.2, 1.0, 0.2]) * max(0, N @ L) + 0.8 * max(0, N @ Hh) ** 100
    assert np.allclose(img["rgba32f"][0, 0, :3], col, atol=2e-5)
    assert img["rgba32f"][0, 0, 3] == 1.0
    # back-facing: ambient only (shader.rgen:104-105)
    cam = rt.CameraVectors.make((0, 0, -20), (-1, 0, 0), (0, 1, 0), (0, 0, 1))
    back = sc.render(cam, p, want=("rgba32f",))["rgba32f"][0, 0, :3]
    assert np.allclose(back, [0.08, 0.24, 0.08], atol=1e-7)


def test_shadow_ray_blocks_light(rt, ob):
    v = np.array([[-50, -50, 0, 0, 0, 1], [50, -50, 0, 0, 0, 1], [0, 50, 0, 0, 0, 1]], np.float32)
    cv, ci = scenes.unit_cube()
    # a cube between the lit point and the light (5,5,5) along the light direction
    blocked = rt.SceneInputs([(v, np.array([[0, 1, 2]], np.uint32)), (cv, ci)],
                             [(0, 0, rt.IDENTITY34), (1, 0, rt.translation34(2.5, 2.5, 2.5))], None)
    p = rt.FrameParams.make(101, 101, 0, 1)   # centre pixel: hit point within 0.1 of the origin
    img = ob.OracleScene(blocked).render(rt.CameraVectors.default(), p, crop=(50, 50, 51, 51), want=("rgba32f", "counts"))
    assert img["counts"]["shadow"] == 1
    assert np.allclose(img["rgba32f"][0, 0, :3], [0.08, 0.24, 0.08], atol=1e-7)


def test_oracle_golden_hashes(rt, ob):
    pins = json.load(open(os.path.join(scenes.GOLDEN, "oracle_hashes.json")))
    cam = rt.CameraVectors.default()
    for name, inputs, params in scenes.hash_cases(rt):
        img = ob.OracleScene(inputs).render(cam, params)["rgba8"]
        assert hashlib.sha256(img.tobytes()).hexdigest() == pins[name], name


def test_invert34_contract(ob):
    m = np.array([0.5, 0, 0.8660254, 1, 0, 2, 0, -3, -0.8660254, 0, 0.5, 4], np.float32)
    inv = ob.invert34(m)
    M = np.eye(4); M[:3] = m.reshape(3, 4).astype(np.float64)
    I = np.eye(4); I[:3] = inv.reshape(3, 4).astype(np.float64)
    assert np.allclose(M @ I, np.eye(4), atol=1e-6)
    assert np.array_equal(ob.invert34(np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 5], np.float32)),
                          np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, -5], np.float32))
