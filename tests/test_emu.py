"""The product's device functions (rt_build.cuh, rt_trav.cuh, rt_shade.cuh) compiled for the
CPU (tests/emu) and checked against the oracle without a GPU: LBVH build, 8-wide collapse and
compression, traversal, shading and the pixel-slot mapping of the wavefront pipeline."""
import numpy as np
import pytest

import scenes


def compare(rt, ob, emu, inputs, params, accel, cam=None, materials=None):
    cam = cam or rt.CameraVectors.default()
    es, osc = emu.EmuScene(inputs, accel), ob.OracleScene(inputs)
    if materials is not None:
        es.set_materials(materials)
        osc.set_materials(materials)
    a = es.render(cam, params)
    b = osc.render(cam, params, want=("rgba8", "rgba32f", "primary", "counts"))
    for f in ("instance", "primitive", "t", "u", "v"):
        assert np.array_equal(a["primary"][f], b["primary"][f]), f
    assert np.array_equal(a["rgba32f"], b["rgba32f"])
    assert np.array_equal(a["rgba8"], b["rgba8"])
    for k in ("primary", "secondary", "shadow", "miss", "hit"):
        assert a["counts"][k] == b["counts"][k], k
    return a["counts"]


@pytest.mark.parametrize("accel", [0, 2, 3, 4])
def test_smoke_scene_bit_exact(rt, ob, emu, accel):
    compare(rt, ob, emu, scenes.smoke_scene(rt), rt.FrameParams.make(96, 72, 6, 3), accel)


def test_materials_bit_exact_and_defaults_are_the_reference_constants(rt, ob, emu):
    """rt_scene_set_materials: emulated kernels == oracle with per-instance Ka/Kd/Ks/Ns/Ni; the default
    material is bit-identical to no material at all; other materials change the picture."""
    inputs = scenes.smoke_scene(rt)
    p = rt.FrameParams.make(96, 72, 6, 3)
    cam = rt.CameraVectors.default()
    inputs.instances[0] = (inputs.instances[0][0], rt.TYPE_DIFFUSE, inputs.instances[0][2])   # more lit surface
    compare(rt, ob, emu, inputs, p, 0, materials=scenes.materials_for(rt, 3))
    es = emu.EmuScene(inputs, 0)
    plain = es.render(cam, p)["rgba32f"]
    es.set_materials([rt.Material.make()] * 3)
    assert np.array_equal(es.render(cam, p)["rgba32f"], plain)
    es.set_materials(scenes.materials_for(rt, 3))
    assert not np.array_equal(es.render(cam, p)["rgba32f"], plain)
    es.set_materials(None)
    assert np.array_equal(es.render(cam, p)["rgba32f"], plain)


def test_det_powf_against_double_pow(ob):
    """the fixed-sequence pow of the material exponent: within 1 ulp of the correctly rounded value"""
    rng = np.random.default_rng(5)
    x = np.concatenate([rng.uniform(0, 1, 4000), [0.0, 1.0, 1e-30, 0.999999, 0.5]]).astype(np.float32)
    y = np.concatenate([rng.uniform(0.5, 900, 4000), [17.5, 100.0, 3.0, 423.0, 1.0]]).astype(np.float32)
    got = np.array([ob.det_powf(a, b) for a, b in zip(x, y)], np.float32)
    ref = np.power(x.astype(np.float64), y.astype(np.float64))
    ulp = np.spacing(np.maximum(ref, 1e-38).astype(np.float32)).astype(np.float64)
    assert np.all(np.abs(got.astype(np.float64) - ref) <= ulp)
    assert ob.det_powf(0.0, 5.0) == 0.0 and ob.det_powf(1.0, 5.0) == 1.0 and ob.det_powf(0.5, 1.0) == 0.5


def test_standin_mesh_refraction_depth8(rt, ob, emu):
    c = compare(rt, ob, emu, scenes.c3_scene(rt, level=4, sky_size=64), rt.FrameParams.make(160, 90, 8, 4), 0)
    assert c["shadow"] > 0 and c["secondary"] > c["primary"] // 20


def test_rotated_instances_and_camera(rt, ob, emu):
    inputs = scenes.smoke_scene(rt)
    xf = rt.orbit_transforms(0.61)
    inputs.instances[0] = (0, rt.TYPE_REFRACTIVE, tuple(xf[1]))
    cam = rt.Camera()
    cam.processMouseMovement(-0.2, 0.1)
    cam.move(rt.LEFT, 2.0)
    compare(rt, ob, emu, inputs, rt.FrameParams.make(90, 60, 5, 2), 0, cam.vectors())


@pytest.mark.parametrize("accel", [0, 3, 4])
def test_many_instances_use_the_top_level_tree(rt, ob, emu, accel):
    compare(rt, ob, emu, scenes.instanced_grid_scene(rt), rt.FrameParams.make(120, 80, 5, 2), accel)


def test_tiny_and_degenerate_meshes(rt, ob, emu):
    cv, ci = scenes.unit_cube()
    one = (np.array([[-3, -3, 0, 0, 0, 1], [3, -3, 0, 0, 0, 1], [0, 3, 0, 0, 0, 1]], np.float32),
           np.array([[0, 1, 2]], np.uint32))
    for accel in (0, 2, 3, 4):
        compare(rt, ob, emu, rt.SceneInputs([one], [(0, 0, rt.IDENTITY34)], None), rt.FrameParams.make(40, 30, 0, 1), accel)
    two = (cv, ci[:2])
    dup = (np.concatenate([cv, cv[:3]]), np.concatenate([ci, ci[:4], np.array([[0, 0, 1], [2, 2, 2]], np.uint32)]))
    empty = (np.zeros((0, 6), np.float32), np.zeros((0, 3), np.uint32))
    inputs = rt.SceneInputs([two, dup, empty], [(0, 1, rt.translation34(-3, 0, 0)), (1, 2, rt.IDENTITY34),
                                                (2, 0, rt.IDENTITY34), (1, 7, rt.translation34(3, 0, 2))],
                            scenes.sea_sky_64())
    for accel in (0, 3, 4):
        compare(rt, ob, emu, inputs, rt.FrameParams.make(80, 60, 4, 2), accel)


def decode_wide(nodes):
    """-> list of (box lo, box hi, children) with children = (slot, kind, lo, hi, payload)."""
    out = []
    for w in nodes:
        p = w[:3].copy().view(np.float32).astype(np.float64)
        e = [(int(w[3]) >> s) & 0xFF for s in (0, 8, 16)]
        imask = int(w[3]) >> 24
        scale = np.array([2.0 ** (x - 127) for x in e])
        child_base, tri_base = int(w[4]), int(w[5])
        by = lambda a, b, i: (int(w[a if i < 4 else b]) >> (8 * (i & 3))) & 0xFF
        kids, rank = [], 0
        for s in range(8):
            meta = by(6, 7, s)
            if meta == 0:
                continue
            lo = p + scale * np.array([by(8, 9, s), by(10, 11, s), by(12, 13, s)])
            hi = p + scale * np.array([by(14, 15, s), by(16, 17, s), by(18, 19, s)])
            if imask & (1 << s):
                assert meta == (0x20 | (24 + s))
                kids.append((s, "node", lo, hi, child_base + rank))
                rank += 1
            else:
                cnt = {1: 1, 3: 2, 7: 3}[meta >> 5]
                kids.append((s, "leaf", lo, hi, (tri_base + (meta & 31), cnt)))
        out.append(kids)
    return out


def test_wide_bvh_structure(rt, emu):
    """Every triangle sits in exactly one leaf, every node is referenced exactly once, and each
    quantised child box contains the geometry below it."""
    v, i = rt.make_standin_armadillo(3)
    es = emu.EmuScene(rt.SceneInputs([(v, i)], [(0, 0, rt.IDENTITY34)], None), rt.ACCEL_WIDE8)
    nodes = es.wide_nodes(0)
    tris = es.leaf_tris(0, i.shape[0])
    prims = tris[:, 3].copy().view(np.uint32)
    assert sorted(prims.tolist()) == list(range(i.shape[0]))
    for k in range(i.shape[0]):
        assert np.array_equal(tris[k, [0, 1, 2, 4, 5, 6, 8, 9, 10]].reshape(3, 3), v[i[prims[k]], :3])
    kids = decode_wide(nodes)
    seen_nodes, seen_tris = {0}, set()

    def bounds(n):
        lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
        for (_, kind, clo, chi, payload) in kids[n]:
            if kind == "node":
                assert payload not in seen_nodes
                seen_nodes.add(payload)
                blo, bhi = bounds(payload)
            else:
                first, cnt = payload
                assert 1 <= cnt <= 3
                ids = list(range(first, first + cnt))
                assert not (seen_tris & set(ids))
                seen_tris.update(ids)
                pts = tris[ids][:, [0, 1, 2, 4, 5, 6, 8, 9, 10]].reshape(-1, 3)
                blo, bhi = pts.min(0), pts.max(0)
            assert np.all(clo <= blo + 1e-9) and np.all(chi >= bhi - 1e-9), "child box does not contain its geometry"
            lo, hi = np.minimum(lo, blo), np.maximum(hi, bhi)
        return lo, hi

    bounds(0)
    assert len(seen_nodes) == nodes.shape[0] and len(seen_tris) == i.shape[0]


@pytest.mark.parametrize("passes", ["0", "2"])
def test_binary_bvh_structure(rt, emu, passes, monkeypatch):
    """After the SAH treelet refinement the binary tree is still a tree: every node is reached
    exactly once from the root, every triangle sits in exactly one leaf and every child box
    contains the geometry below it.  Refinement must lower the SAH cost."""
    monkeypatch.setenv("RT_TREELET_PASSES", passes)
    v, i = rt.make_standin_armadillo(3)
    es = emu.EmuScene(rt.SceneInputs([(v, i)], [(0, 0, rt.IDENTITY34)], None), rt.ACCEL_BVH2)
    nodes = es.bvh2_nodes(0)
    ntri = i.shape[0]
    assert nodes.shape[0] == ntri - 1
    tris = es.leaf_tris(0, ntri)
    prims = tris[:, 3].copy().view(np.uint32)
    assert sorted(prims.tolist()) == list(range(ntri))
    child = nodes[:, 12:14].copy().view(np.int32)
    seen_nodes, seen_leaves = {0}, set()
    area = [0.0]

    def bounds(n):
        lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
        boxes = [(nodes[n, 0:3], nodes[n, 3:6]), (nodes[n, 6:9], nodes[n, 9:12])]
        for c, (clo, chi) in zip(child[n], boxes):
            if c >= 0:
                assert c not in seen_nodes
                seen_nodes.add(int(c))
                blo, bhi = bounds(int(c))
            else:
                slot = ~int(c)
                assert slot not in seen_leaves
                seen_leaves.add(slot)
                pts = tris[slot][[0, 1, 2, 4, 5, 6, 8, 9, 10]].reshape(3, 3)
                blo, bhi = pts.min(0), pts.max(0)
            assert np.all(clo <= blo + 1e-9) and np.all(chi >= bhi - 1e-9), "child box does not contain its geometry"
            d = chi - clo
            area[0] += float(d[0] * d[1] + d[1] * d[2] + d[2] * d[0])
            lo, hi = np.minimum(lo, clo), np.maximum(hi, chi)
        return lo, hi

    import sys
    sys.setrecursionlimit(10000)
    bounds(0)
    assert len(seen_nodes) == ntri - 1 and len(seen_leaves) == ntri
    test_binary_bvh_structure.area[passes] = area[0]
    if len(test_binary_bvh_structure.area) == 2:
        assert test_binary_bvh_structure.area["2"] < 0.97 * test_binary_bvh_structure.area["0"]


test_binary_bvh_structure.area = {}


def test_treelet_dp_finds_the_optimum(emu):
    """The treelet optimiser (dynamic programme over leaf subsets, splits read in groups of one, three and
    four) against an exhaustive search in double precision: for 2..7 random leaf boxes the cost it reports is
    the minimum over all binary topologies (or the unchanged chain when the gain is below 0.01 %), and the
    tree it writes is a tree over exactly those leaves with that cost."""
    from functools import lru_cache
    rng = np.random.default_rng(5)
    CT, CI = 1.0, 1.3

    def half_area(lo, hi):
        d = hi - lo
        return float(d[0] * d[1] + d[1] * d[2] + d[2] * d[0])

    for trial in range(60):
        n = 2 + trial % 6
        lo = rng.uniform(-4, 4, (n, 3))
        hi = lo + rng.uniform(0.01, 2.5, (n, 3)) * rng.choice([0.2, 1.0, 3.0])
        boxes = np.concatenate([lo, hi], 1).astype(np.float32)
        lo, hi = boxes[:, :3].astype(np.float64), boxes[:, 3:].astype(np.float64)
        # leaf costs: single triangles (CI * area) or whole subtrees (larger)
        lcost = np.array([CI * half_area(lo[j], hi[j]) * rng.choice([1.0, 1.0, 4.0]) for j in range(n)], np.float32)

        def area(mask):
            idx = [j for j in range(n) if mask >> j & 1]
            return half_area(lo[idx].min(0), hi[idx].max(0))

        @lru_cache(None)
        def best(mask):
            if mask & (mask - 1) == 0:
                return float(lcost[mask.bit_length() - 1])
            sub, m = (mask - 1) & mask, float("inf")
            while sub:
                if sub < (mask ^ sub):          # each unordered split once
                    m = min(m, best(sub) + best(mask ^ sub))
                sub = (sub - 1) & mask
            return CT * area(mask) + m

        before, after, left, right = emu.treelet(boxes, lcost)
        opt = best((1 << n) - 1)
        chain = sum(CT * area(((1 << n) - 1) & ~((1 << i) - 1)) for i in range(n - 1)) + float(lcost.sum())
        assert before == pytest.approx(chain, rel=1e-5)
        assert opt <= chain * (1 + 1e-7)
        if opt < 0.9998 * chain:
            assert after == pytest.approx(opt, rel=2e-5), (n, trial)
        elif opt > 0.99995 * chain:
            assert after == pytest.approx(chain, rel=2e-5)
        else:
            assert opt * (1 - 2e-5) <= after <= chain * (1 + 2e-5)
        # the tree that was written: every leaf once, its SAH cost is the reported one
        L = n - 1
        seen = []

        def walk(node):
            if node >= L:
                seen.append(node - L)
                return 1 << (node - L), float(lcost[node - L])
            ml, cl = walk(int(left[node]))
            mr, cr = walk(int(right[node]))
            return ml | mr, CT * area(ml | mr) + cl + cr

        mask, cost = walk(0)
        assert sorted(seen) == list(range(n)) and mask == (1 << n) - 1
        assert cost == pytest.approx(after, rel=2e-5)


def test_trace_rays_matches_oracle_brute_force(rt, ob, emu):
    inputs = scenes.c3_scene(rt, level=3, sky_size=8)
    rng = np.random.default_rng(11)
    n = 6000
    o = rng.normal(0, 6, (n, 3)).astype(np.float32)
    d = (rng.normal(0, 1.5, (n, 3)) + [0, 1, 2.5] - o).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    es, osc = emu.EmuScene(inputs, 0), ob.OracleScene(inputs)
    for tmax in (10000.0, 7.0):
        a, b = es.trace_rays(o, d, 0.001, tmax), osc.trace_rays(o, d, 0.001, tmax, brute=True)
        for f in ("instance", "primitive", "t", "u", "v"):
            assert np.array_equal(a[f], b[f])
        a, b = es.trace_rays(o, d, 0.001, tmax, True), osc.trace_rays(o, d, 0.001, tmax, True, brute=True)
        assert np.array_equal(a["instance"], b["instance"])


@pytest.mark.parametrize("w,h,tw,th,n", [(200, 136, 32, 16, 3), (64, 64, 64, 64, 1), (70, 33, 8, 4, 4)])
def test_slot_to_pixel_is_a_bijection(rt, emu, w, h, tw, th, n):
    """Every pixel of the frame is owned by exactly one (shard, slot); padding slots map outside."""
    import ctypes as C
    seen = np.zeros((h, w), np.int32)
    for s in range(n):
        p = rt.FrameParams.make(w, h, 1, 1, tile_w=tw, tile_h=th, shard_index=s, shard_count=n)
        tiles = ((w + tw - 1) // tw) * ((h + th - 1) // th)
        local = (tiles - s + n - 1) // n if tiles > s else 0
        for slot in range(local * tw * th):
            ok, x, y = emu.slot_to_pixel(p, slot)
            if ok:
                seen[y, x] += 1
                assert ((y // th) * ((w + tw - 1) // tw) + x // tw) % n == s
    assert (seen == 1).all()


def test_node_tests_are_conservative(emu):
    """The box test only has to be conservative: a ray that reaches a point of a mesh must pass the
    test of every padded box around that point when the far limit is exactly the distance of the point (best.t == t of
    the hit, the tightest the traversal ever makes it).  Checked for the round-1 test and the two round-2 tests (one fma
    per plane: the default; near / far planes picked by sign) against float64 geometry: origins from inside the mesh out
    to 10 000 mesh extents, axis-parallel and near-parallel directions, flat and point-sized boxes, points on box faces."""
    rng = np.random.default_rng(5)
    n = 400000
    M = 10.0 ** rng.uniform(-2, 3, n)                                   # max |coordinate| of the mesh the box belongs to
    c = rng.uniform(-1, 1, (n, 3)) * M[:, None] * 0.9                  # box centre inside the mesh bounds
    half = 10.0 ** rng.uniform(-6, -0.5, (n, 3)) * M[:, None]
    half[rng.random((n, 3)) < 0.15] = 0.0                              # flat / degenerate boxes
    lo, hi = c - half, c + half
    p = lo + rng.random((n, 3)) * (hi - lo)                            # the point the ray reaches ...
    face = rng.random(n) < 0.3                                         # ... sometimes exactly on a face of the box
    ax = rng.integers(0, 3, n)
    p[face, ax[face]] = np.where(rng.random(face.sum()) < 0.5, lo[face, ax[face]], hi[face, ax[face]])
    d = rng.normal(0, 1, (n, 3))
    par = rng.random(n) < 0.3                                          # axis-parallel and nearly parallel directions
    k = rng.integers(0, 3, n)
    for a in range(3):
        sel = par & (k != a)
        d[sel, a] *= 10.0 ** rng.uniform(-30, -3, sel.sum())
    d[rng.random(n) < 0.05, rng.integers(0, 3)] = 0.0
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    dist = 10.0 ** rng.uniform(-3, 4, n) * M                           # origin 0.001 ... 10 000 mesh extents away
    o64 = p - d * dist[:, None]
    # what the kernels see: float32 box (leaf padding of rt_build.cu: 4e-6 * max |coordinate|, rounded outward), float32 ray
    pad = (M * 4e-6 + 1e-30)[:, None]
    lo32 = np.nextafter((lo - pad).astype(np.float32), np.float32(-np.inf))
    hi32 = np.nextafter((hi + pad).astype(np.float32), np.float32(np.inf))
    o32, d32 = o64.astype(np.float32), d.astype(np.float32)
    ok = (np.linalg.norm(d32, axis=1) > 0.5)
    # distance of the point along the float32 ray, as the triangle test would report it (float64, rounded up one ulp)
    dd = d32.astype(np.float64)
    t = ((p - o32.astype(np.float64)) * dd).sum(1) / (dd * dd).sum(1)
    # the float32 ray does not pass exactly through p any more: keep the pairs whose ray still comes within the padding
    closest = o32.astype(np.float64) + dd * t[:, None]
    ok &= (np.abs(closest - p) <= 0.25 * pad).all(1) & (t > 0.002) & (t < 9000.0)
    tmax = np.nextafter(t.astype(np.float32), np.float32(np.inf))
    m = emu.slab_tests(o32[ok], d32[ok], np.concatenate([lo32, hi32], 1)[ok], 0.001, tmax[ok])
    assert ok.sum() > 100000
    for bit, name in ((1, "slab_interval"), (2, "slab_interval_fma"), (4, "slab_interval_sel")):
        missed = np.flatnonzero((m & bit) == 0)
        assert missed.size == 0, (name, missed.size, missed[:5])
