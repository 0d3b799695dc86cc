"""Parity of the CUDA path (through the C ABI) with the CPU oracle.  Integer results
(instance / primitive ids, ray counts) must match exactly; radiance follows the shared FP32
numeric contract, so RGBA32F is expected bit-identical and RGBA8 identical; the stated
tolerance from BASELINE.json (<= 2/255 per channel, PSNR >= 45 dB, >= 99.99 % primary ids) is
asserted as the floor."""
import os

import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu


def psnr(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


def check_frame(rt, ob, inputs, params, accel=0, cam=None, exact=True):
    cam = cam or rt.CameraVectors.default()
    scene = rt.Scene(inputs, device_id=0, accel=accel)
    p32 = params.copy()
    p32.output_format = rt.FORMAT_RGBA32F
    img8 = scene.render(cam, params)
    st = scene.stats()
    img32 = scene.render(cam, p32)
    prim = scene.primary_hits(cam, params)
    ref = ob.OracleScene(inputs).render(cam, params, want=("rgba8", "rgba32f", "primary", "counts"))
    same = (prim["instance"] == ref["primary"]["instance"]) & (prim["primitive"] == ref["primary"]["primitive"])
    assert same.mean() >= 0.9999
    diff = np.abs(img8.astype(np.int32) - ref["rgba8"].astype(np.int32))
    assert diff.max() <= 2, f"max per-channel error {diff.max()}/255"
    assert psnr(img8, ref["rgba8"]) >= 45.0
    c = ref["counts"]
    if exact:
        assert same.all()
        assert np.array_equal(prim["t"], ref["primary"]["t"]) and np.array_equal(prim["u"], ref["primary"]["u"])
        assert np.array_equal(img32, ref["rgba32f"])
        assert np.array_equal(img8, ref["rgba8"])
        assert (st["rays_primary"], st["rays_secondary"], st["rays_shadow"], st["n_miss"], st["n_hit"]) == \
               (c["primary"], c["secondary"], c["shadow"], c["miss"], c["hit"])
    scene.close()
    return st


@pytest.mark.parametrize("accel", [0, 2, 3, 4])
def test_smoke_scene_all_accels(gpu_rt, ob, accel):
    check_frame(gpu_rt, ob, scenes.smoke_scene(gpu_rt), gpu_rt.FrameParams.make(160, 120, 6, 4), accel)


def test_golden_hashes_through_cuda(gpu_rt):
    """The CUDA frames must hash to the committed oracle pins (tests/golden/oracle_hashes.json)."""
    import hashlib, json, os
    pins = json.load(open(os.path.join(scenes.GOLDEN, "oracle_hashes.json")))
    for name, inputs, params in scenes.hash_cases(gpu_rt):
        sc = gpu_rt.Scene(inputs, device_id=0)
        img = sc.render(gpu_rt.CameraVectors.default(), params)
        assert hashlib.sha256(img.tobytes()).hexdigest() == pins[name], name
        sc.close()


def test_c2_teapots_depth4(gpu_rt, ob):
    check_frame(gpu_rt, ob, scenes.c2_scene(gpu_rt, 256), gpu_rt.FrameParams.make(480, 270, 4, 4))


def test_c3_armadillo_standin_depth8(gpu_rt, ob):
    check_frame(gpu_rt, ob, scenes.c3_scene(gpu_rt, level=5, sky_size=256), gpu_rt.FrameParams.make(384, 216, 8, 4))


def test_materials_match_the_oracle_and_default_to_the_shader_constants(gpu_rt, ob):
    """rt_scene_set_materials: per-instance Ka/Kd/Ks/Ns/Ni, bit-exact against the
    oracle; the default material and NULL both reproduce the reference's constants."""
    rt = gpu_rt
    inputs = scenes.c3_scene(rt, level=4, sky_size=128)
    cam = rt.CameraVectors.default()
    p = rt.FrameParams.make(320, 180, 8, 4, output_format=rt.FORMAT_RGBA32F)
    sc = rt.Scene(inputs, device_id=0)
    plain = sc.render(cam, p)
    mats = scenes.materials_for(rt, len(inputs.instances))
    sc.set_materials(mats)
    got = sc.render(cam, p)
    osc = ob.OracleScene(inputs)
    osc.set_materials(mats)
    ref = osc.render(cam, p, want=("rgba32f",))["rgba32f"]
    assert np.array_equal(got, ref)
    assert not np.array_equal(got, plain)
    sc.set_materials([rt.Material.make()] * len(inputs.instances))
    assert np.array_equal(sc.render(cam, p), plain)
    sc.set_materials(mats)
    sc.set_materials(None)
    assert np.array_equal(sc.render(cam, p), plain)
    with pytest.raises(rt.RtError):
        sc.set_materials(mats[:1] if len(mats) > 1 else mats * 2)
    sc.close()


def test_default_config_depth63_odd_size(gpu_rt, ob):
    """config.h defaults (63 bounces, 4 spp) on a frame that is not a multiple of the tile size."""
    check_frame(gpu_rt, ob, scenes.c1_like_scene(gpu_rt), gpu_rt.FrameParams.make(203, 151, 63, 4))


def test_moved_camera_and_orbit_transform(gpu_rt, ob):
    rt = gpu_rt
    cam = rt.Camera()
    cam.processMouseMovement(0.35, -0.12)
    cam.move(rt.RIGHT, 3.0)
    cam.move(rt.UP, 1.5)
    inputs = scenes.smoke_scene(rt)
    xf = rt.orbit_transforms(0.37)
    inputs.instances[0] = (inputs.instances[0][0], inputs.instances[0][1], tuple(xf[0]))
    inputs.instances[1] = (inputs.instances[1][0], inputs.instances[1][1], tuple(xf[1]))
    check_frame(rt, ob, inputs, rt.FrameParams.make(200, 120, 8, 2), cam=cam.vectors())


def test_set_transforms_matches_rebuild(gpu_rt, ob):
    rt = gpu_rt
    inputs = scenes.smoke_scene(rt)
    sc = rt.Scene(inputs, device_id=0)
    p = rt.FrameParams.make(160, 96, 5, 2)
    cam = rt.CameraVectors.default()
    xf = np.array([i[2] for i in inputs.instances], np.float32)
    xf[:2] = rt.orbit_transforms(1.3)
    sc.set_transforms(xf)
    img = sc.render(cam, p)
    osc = ob.OracleScene(inputs)
    osc.set_transforms(xf)
    assert np.array_equal(img, osc.render(cam, p)["rgba8"])
    sc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("accel", [0, 3, 4])
def test_instanced_grid_top_level_tree(gpu_rt, ob, accel):
    """13 instances (> RT_TLAS_MIN_INSTANCES): rays walk the top-level tree; also after a transform update."""
    rt = gpu_rt
    inputs = scenes.instanced_grid_scene(rt)
    check_frame(rt, ob, inputs, rt.FrameParams.make(240, 160, 5, 4), accel)


def test_headless_host_frames_match_the_python_binding(gpu_rt, tmp_path):
    """rt_headless renders the orbit animation through the C ABI from C++; its PPM frames are
    bit-identical to the same frames rendered through the ctypes binding."""
    import subprocess
    rt = gpu_rt
    exe = os.path.join(rt.REPO_ROOT, "vulkan-raytracing_b200", "rt_headless")
    prefix = str(tmp_path / "f")
    r = subprocess.run([exe, "--center", "standin:3", "--orbiting", "standin:4", "--center-type", "1", "--orbiting-type", "2",
                        "--skybox", "procedural:64", "--size", "160x120", "--bounces", "5", "--spp", "2", "--frames", "3",
                        "--dt", "0.5", "--out", prefix, "--fps"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "FPS:" in r.stdout
    a, ai = rt.make_standin_armadillo(3)
    b, bi = rt.make_standin_armadillo(4)
    inputs = rt.SceneInputs([(a, ai), (b, bi)], [(0, rt.TYPE_MIRROR, rt.IDENTITY34), (1, rt.TYPE_REFRACTIVE, rt.IDENTITY34)],
                            rt.make_procedural_skybox(64))
    scene = rt.Scene(inputs, device_id=0)
    p = rt.FrameParams.make(160, 120, 5, 2)
    cam = rt.Camera().vectors()
    for f in range(3):
        scene.set_transforms(rt.orbit_transforms(np.float32(f) * np.float32(0.5) * np.float32(0.1)))
        img = scene.render(cam, p)
        with open(f"{prefix}_{f:04d}.ppm", "rb") as fh:
            assert fh.readline().strip() == b"P6"
            w, h = map(int, fh.readline().split())
            assert fh.readline().strip() == b"255"
            ppm = np.frombuffer(fh.read(), np.uint8).reshape(h, w, 3)
        assert np.array_equal(ppm, img[:, :, :3]), f"frame {f} differs"
    scene.close()


def test_trace_rays_closest_and_any(gpu_rt, ob):
    rt = gpu_rt
    inputs = scenes.c3_scene(rt, level=4, sky_size=16)
    rng = np.random.default_rng(7)
    n = 20000
    o = rng.normal(0, 6, (n, 3)).astype(np.float32)
    tgt = rng.normal(0, 1.5, (n, 3)).astype(np.float32) + np.array([0, 1, 2.5], np.float32)
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    sc = rt.Scene(inputs, device_id=0)
    osc = ob.OracleScene(inputs)
    for tmax in (10000.0, 6.0):
        a, b = sc.trace_rays(o, d, 0.001, tmax), osc.trace_rays(o, d, 0.001, tmax, brute=True)
        for f in ("instance", "primitive", "t", "u", "v"):
            assert np.array_equal(a[f], b[f]), f
        a, b = sc.trace_rays(o, d, 0.001, tmax, any_hit=True), osc.trace_rays(o, d, 0.001, tmax, any_hit=True, brute=True)
        assert np.array_equal(a["instance"], b["instance"])
    assert (a["instance"] != rt.MISS).mean() > 0.2
    sc.close()


@pytest.mark.parametrize("accel", [0, 3, 4])
def test_axis_parallel_and_far_rays(gpu_rt, ob, accel):
    """Box culling must stay conservative where the slab arithmetic degenerates: directions with
    exactly-zero components (1/d clamped), origins lying on box planes, origins far from the
    scene.  Closest hits must equal the oracle's brute force bit for bit."""
    rt = gpu_rt
    inputs = scenes.c3_scene(rt, level=3, sky_size=8)
    rng = np.random.default_rng(11)
    n = 6000
    axes = np.eye(3, dtype=np.float32)
    d = np.concatenate([axes[rng.integers(0, 3, n)] * rng.choice([-1.0, 1.0], (n, 1)).astype(np.float32),      # axis parallel
                        np.stack([rng.normal(size=n), np.zeros(n), rng.normal(size=n)], 1).astype(np.float32),  # one zero component
                        rng.normal(size=(n, 3)).astype(np.float32)])
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    tgt = rng.normal(0, 1.2, (3 * n, 3)).astype(np.float32) + np.array([0, 0.5, 2.5], np.float32)
    dist = np.concatenate([rng.uniform(3, 12, 2 * n), rng.uniform(2000, 9000, n)]).astype(np.float32)[:, None]
    o = (tgt - d * dist).astype(np.float32)
    o[:n:3] = np.round(o[:n:3], 0)          # many origins with integer coordinates (on box planes of the cube-ish teapot lid)
    sc = rt.Scene(inputs, device_id=0, accel=accel)
    osc = ob.OracleScene(inputs)
    a, b = sc.trace_rays(o, d, 0.001, 10000.0), osc.trace_rays(o, d, 0.001, 10000.0, brute=True)
    for f in ("instance", "primitive", "t", "u", "v"):
        assert np.array_equal(a[f], b[f]), f
    assert (a["instance"] != rt.MISS).mean() > 0.1
    with pytest.raises(rt.RtError):
        sc.trace_rays(o[:4], d[:4], -1.0, 10.0)   # traceRayEXT requires tmin >= 0
    sc.close()


def test_edge_cases(gpu_rt, ob):
    rt = gpu_rt
    cv, ci = scenes.unit_cube()
    cam = rt.CameraVectors.default()
    # single triangle mesh, 1 spp, 0 bounces, no skybox
    tri_v = np.array([[-3, -3, 0, 0, 0, 1], [3, -3, 0, 0, 0, 1], [0, 3, 0, 0, 0, 1]], np.float32)
    one = rt.SceneInputs([(tri_v, np.array([[0, 1, 2]], np.uint32))], [(0, rt.TYPE_DIFFUSE, rt.IDENTITY34)], None)
    check_frame(rt, ob, one, rt.FrameParams.make(64, 48, 0, 1))
    # an empty mesh next to a cube; unknown material type 7 (the reference's chain falls through)
    empty = (np.zeros((0, 6), np.float32), np.zeros((0, 3), np.uint32))
    mix = rt.SceneInputs([empty, (cv, ci)], [(0, rt.TYPE_MIRROR, rt.IDENTITY34), (1, 7, rt.translation34(0, 0, 8)),
                                             (1, rt.TYPE_REFRACTIVE, rt.translation34(2.5, 0, 8))], scenes.sea_sky_64())
    check_frame(rt, ob, mix, rt.FrameParams.make(96, 64, 3, 3))
    # degenerate (zero-area) and duplicated triangles
    v = np.concatenate([cv, cv[:3]])
    i = np.concatenate([ci, ci[:4], np.array([[0, 0, 1], [2, 2, 2]], np.uint32)])
    check_frame(rt, ob, rt.SceneInputs([(v, i)], [(0, rt.TYPE_REFRACTIVE, rt.IDENTITY34)], scenes.sea_sky_64()),
                rt.FrameParams.make(80, 80, 6, 2))


def test_error_paths(gpu_rt):
    rt = gpu_rt
    cv, ci = scenes.unit_cube()
    bad = ci.copy()
    bad[0, 0] = 999
    with pytest.raises(rt.RtError):
        rt.Scene(rt.SceneInputs([(cv, bad)], [(0, 0, rt.IDENTITY34)], None), device_id=0)
    sc = rt.Scene(rt.SceneInputs([(cv, ci)], [(0, 0, rt.IDENTITY34)], None), device_id=0)
    with pytest.raises(rt.RtError):
        sc.render(rt.CameraVectors.default(), rt.FrameParams.make(0, 10, 1, 1))
    with pytest.raises(rt.RtError):
        sc.render(rt.CameraVectors.default(), rt.FrameParams.make(16, 16, 1, 1, tile_w=12))
    sc.close()


def test_device_normals_equal_the_host_loader(gpu_rt):
    """rt_generate_normals == rth_generate_normals bit for bit: full-size stand-in mesh,
    an unused vertex, a degenerate triangle, a vertex shared by very many triangles, an empty mesh."""
    rt = gpu_rt
    v, i = rt.make_standin_armadillo(7)
    rng = np.random.default_rng(3)
    v = v.copy()
    v[:, :3] += rng.normal(0, 1e-3, (v.shape[0], 3)).astype(np.float32)
    v[:, 3:] = 0
    v = np.vstack([v, np.array([[9, 9, 9, 0, 0, 0]], np.float32)])
    fan = np.stack([np.zeros(5000, np.uint32), rng.integers(1, 4000, 5000).astype(np.uint32),
                    rng.integers(1, 4000, 5000).astype(np.uint32)], 1)
    i = np.vstack([i, np.array([[0, 0, 1]], np.uint32), fan])
    dev, ms = rt.generate_normals(v, i)
    host = rt.generate_normals_host(v, i)
    assert np.array_equal(dev, host)
    assert ms > 0
    # RT_FLAG_GENERATE_NORMALS: a scene built from the normal-less mesh renders like one built from host normals
    sky = rt.make_procedural_skybox(32)
    small_v, small_i = rt.make_standin_armadillo(3)
    bare = small_v.copy()
    bare[:, 3:] = 0
    cam, p = rt.CameraVectors.default(), rt.FrameParams.make(160, 120, 3, 2, output_format=rt.FORMAT_RGBA32F)
    inst = [(0, rt.TYPE_DIFFUSE, rt.IDENTITY34)]
    a = rt.Scene(rt.SceneInputs([(bare, small_i)], inst, sky), device_id=0, accel=rt.FLAG_GENERATE_NORMALS)
    b = rt.Scene(rt.SceneInputs([(rt.generate_normals_host(bare, small_i), small_i)], inst, sky), device_id=0)
    assert np.array_equal(a.render(cam, p), b.render(cam, p))
    a.close(); b.close()
    e, _ = rt.generate_normals(np.zeros((3, 6), np.float32), np.zeros((0, 3), np.uint32))
    assert np.array_equal(e[:, 3:], [[0, 1, 0]] * 3)
    with pytest.raises(rt.RtError):
        rt.generate_normals(np.zeros((3, 6), np.float32), np.array([[0, 1, 3]], np.uint32))


def test_sort_pairs_matches_numpy(gpu_rt):
    rng = np.random.default_rng(3)
    for n in (1, 31, 4096, 4097, 300001, 3000001):   # the last one: more tiles than resident blocks
        k = rng.integers(0, 2 ** 63, n, dtype=np.uint64)
        k[::7] = k[0]  # duplicates: stability matters
        v = np.arange(n, dtype=np.uint32)
        ko, vo = gpu_rt.sort_pairs(k, v)
        order = np.argsort(k, kind="stable")
        assert np.array_equal(ko, k[order]) and np.array_equal(vo, v[order])


def test_sharded_frame_is_bit_identical(gpu_rt):
    """Tile-sharded rendering (multi-GPU path) assembled on one device equals the full frame."""
    import torch
    rt = gpu_rt
    inputs = scenes.smoke_scene(rt)
    sc = rt.Scene(inputs, device_id=0)
    cam = rt.CameraVectors.default()
    for fmt, dt in ((rt.FORMAT_RGBA8, torch.uint8), (rt.FORMAT_RGBA32F, torch.float32)):
        p = rt.FrameParams.make(200, 136, 5, 2, output_format=fmt, tile_w=32, tile_h=16)
        full = sc.render(cam, p)
        n = 3
        q = p.copy()
        q.shard_count, q.output_location = n, rt.LOCATION_DEVICE
        per = int(rt.rt_lib().rt_frame_shard_pixels(q))
        gathered = torch.zeros((n, per, 4), dtype=dt, device="cuda:0")
        for s in range(n):
            q.shard_index = s
            sc.render_into(cam, q, gathered[s].data_ptr())
        frame = torch.zeros((p.height, p.width, 4), dtype=dt, device="cuda:0")
        sc.assemble_shards(q, gathered.data_ptr(), frame.data_ptr())
        assert np.array_equal(frame.cpu().numpy(), full)
    sc.close()


def test_shards_stored_in_frame_equal_the_full_frame(gpu_rt):
    """RT_SHARD_IN_FRAME (the collective-free gather): every shard stores its tiles at their final
    position of one frame; pixels of the other shards stay untouched; a frame created by
    rt_peer_frame_create is an ordinary device buffer for its owner."""
    import torch
    rt = gpu_rt
    inputs = scenes.smoke_scene(rt)
    sc = rt.Scene(inputs, device_id=0)
    cam = rt.CameraVectors.default()
    for fmt, dt, px in ((rt.FORMAT_RGBA8, torch.uint8, 4), (rt.FORMAT_RGBA32F, torch.float32, 16)):
        p = rt.FrameParams.make(200, 136, 5, 2, output_format=fmt, tile_w=32, tile_h=16)
        full = sc.render(cam, p)
        n = 3
        q = p.copy()
        q.shard_count, q.output_location, q.shard_layout = n, rt.LOCATION_DEVICE, rt.SHARD_IN_FRAME
        assert int(rt.rt_lib().rt_frame_output_bytes(q)) == 200 * 136 * px
        pf = rt.PeerFrame.create(0, 200 * 136 * px)
        assert len(pf.handle_bytes()) == 64 and any(pf.handle_bytes())
        frame = torch.as_tensor(pf, device="cuda:0").view(dt).view(136, 200, 4)
        frame.fill_(7)
        seen = np.zeros((136, 200), bool)
        for s in range(n):
            q.shard_index = s
            sc.render_into(cam, q, pf.ptr)
            now = frame.cpu().numpy()
            ty, tx = np.mgrid[0:136, 0:200]
            mine = ((ty // 16) * ((200 + 31) // 32) + tx // 32) % n == s
            seen |= mine
            assert np.array_equal(now[seen], full[seen])
            assert np.all(now[~seen] == 7)
        assert seen.all()
        # arrival flags behind the frame: three ranks report frame 5, the owner's wait for three returns,
        # a wait for a fourth that never reports times out with an error
        for r in range(3):
            pf.signal(r, 5)
        pf.wait(3, 5)
        pf.wait(3, 4)          # an older frame number is already satisfied (wrap-safe compare)
        with pytest.raises(rt.RtError):
            pf.wait(4, 5, timeout_ms=30)
        assert np.array_equal(frame.cpu().numpy(), full)      # the flags are not part of the frame
        del frame
        pf.release()
    # host output of a partial frame is refused
    bad = rt.FrameParams.make(64, 64, 1, 1, shard_count=2, shard_layout=rt.SHARD_IN_FRAME)
    with pytest.raises(rt.RtError):
        sc.render(cam, bad)
    sc.close()


def test_host_frame_receives_the_tiles_of_every_shard(gpu_rt):
    """rt_host_frame_*: the frame lives in shared pinned host memory, k_resolve stores into it through the
    mapped device address; a second mapping of the same name (what another rank's process would open) sees
    the same pixels and flags; whole frame and 3 shards both equal the frame rendered on the device."""
    import os
    rt = gpu_rt
    inputs = scenes.smoke_scene(rt)
    sc = rt.Scene(inputs, device_id=0)
    cam = rt.CameraVectors.default()
    W, H = 200, 136
    p = rt.FrameParams.make(W, H, 5, 2, tile_w=32, tile_h=16)
    full = sc.render(cam, p)
    name = f"/rt_b200_test_{os.getpid()}"
    hf = rt.HostFrame.create(name, 0, W * H * 4)
    try:
        with pytest.raises(rt.RtError):
            rt.HostFrame.create(name, 0, W * H * 4)       # the name is taken
        other = rt.HostFrame.open(name, 0, W * H * 4)       # second mapping = another process's view
        pix = hf.array(H, W, 4)
        q = p.copy()
        q.output_location, q.shard_layout = rt.LOCATION_DEVICE, rt.SHARD_IN_FRAME
        pix[:] = 7
        sc.render_into(cam, q, hf.dev_ptr)                  # one shard = the whole frame
        assert np.array_equal(pix, full) and np.array_equal(other.array(H, W, 4), full)
        pix[:] = 7
        q.shard_count = 3
        for s in (2, 0, 1):
            q.shard_index = s
            sc.render_into(cam, q, other.dev_ptr if s == 1 else hf.dev_ptr)
            hf.signal(s, 9) if s != 1 else other.signal(s, 9)
        hf.wait(3, 9)
        assert np.array_equal(pix, full)
        with pytest.raises(rt.RtError):
            hf.wait(4, 9, timeout_ms=30)
        other.release()
    finally:
        hf.release()
    with pytest.raises(rt.RtError):
        rt.HostFrame.open(name, 0, W * H * 4)               # the creator removed the name
    sc.close()


def test_full_size_properties(gpu_rt, ob):
    """BASELINE C3 size (3840x2160, depth 8): size-independent checks — a centred crop rendered by
    the oracle matches the same pixels of the full CUDA frame exactly, and ray counts are consistent."""
    rt = gpu_rt
    inputs = scenes.c3_scene(rt, level=7, sky_size=512)
    sc = rt.Scene(inputs, device_id=0)
    cam = rt.CameraVectors.default()
    p = rt.FrameParams.make(3840, 2160, 8, 4)
    img = sc.render(cam, p)
    st = sc.stats()
    assert st["rays_primary"] == 3840 * 2160 * 4
    assert st["n_miss"] + st["n_hit"] == st["rays_primary"] + st["rays_secondary"]
    crop = (1800, 1000, 2056, 1096)
    ref = ob.OracleScene(inputs).render(cam, p, crop=crop)["rgba8"]
    assert np.array_equal(img[crop[1]:crop[3], crop[0]:crop[2]], ref)
    sc.close()


# ---- round 2: every BASELINE config at its real size, an independent float64 check, deep trees ------------
def test_c1_exactly_as_baseline_states_it(gpu_rt, ob):
    """Config C1: cube_scene.obj (corrected-normal policy), sea skybox (256x256 fixture of it), 800x600,
    MAX_BOUNCE_COUNT 63, 4 spp: bit-identical to the oracle and to its committed hash."""
    import hashlib, json
    rt = gpu_rt
    st = check_frame(rt, ob, scenes.c1_scene(rt), scenes.c1_params(rt))
    assert st["rays_primary"] == 800 * 600 * 4
    pins = json.load(open(os.path.join(scenes.GOLDEN, "oracle_hashes.json")))
    sc = rt.Scene(scenes.c1_scene(rt), device_id=0)
    img = sc.render(rt.CameraVectors.default(), scenes.c1_params(rt))
    assert hashlib.sha256(img.tobytes()).hexdigest() == pins["c1_cube_scene_800x600_d63_sky256"]
    sc.close()


def test_c2_full_size(gpu_rt, ob):
    """Config C2 at its real size: 1920x1080, depth 4, 2048x2048 skybox faces; whole frame, bit-identical."""
    rt = gpu_rt
    inputs = scenes.c2_scene(rt, 2048)
    p = rt.FrameParams.make(1920, 1080, 4, 4)
    cam = rt.CameraVectors.default()
    sc = rt.Scene(inputs, device_id=0)
    img = sc.render(cam, p)
    st = sc.stats()
    ref = ob.OracleScene(inputs).render(cam, p, want=("rgba8", "counts"))
    assert np.array_equal(img, ref["rgba8"])
    c = ref["counts"]
    assert (st["rays_primary"], st["rays_secondary"], st["rays_shadow"], st["n_miss"], st["n_hit"]) == \
           (c["primary"], c["secondary"], c["shadow"], c["miss"], c["hit"])
    sc.close()


def test_c4_as_eight_shards_full_size(gpu_rt, ob):
    """Config C4 at its real size (7680x4320, depth 8), rendered the way eight GPUs render it: eight interleaved
    tile shards, each stored at its final position of one frame (RT_SHARD_IN_FRAME).  Crops that straddle tile
    boundaries (and so shard boundaries: neighbouring tiles belong to different shards) equal the oracle's pixels."""
    import torch
    rt = gpu_rt
    inputs = scenes.c3_scene(rt, level=7, sky_size=512)
    sc = rt.Scene(inputs, device_id=0)
    cam = rt.CameraVectors.default()
    W, H, n = 7680, 4320, 8
    frame = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda:0")
    p = rt.FrameParams.make(W, H, 8, 4, output_location=rt.LOCATION_DEVICE, shard_count=n, shard_layout=rt.SHARD_IN_FRAME)
    rays = 0
    for s in range(n):
        p.shard_index = s
        sc.render_into(cam, p, frame.data_ptr())
        st = sc.stats()
        rays += st["rays_primary"]
        assert st["stack_overflow"] == 0
    assert rays >= W * H * 4      # (padding tiles at the frame edge add primary slots that are never traced)
    q = rt.FrameParams.make(W, H, 8, 4)
    ref_scene = ob.OracleScene(inputs)
    # 64x64 tiles: x = 3776..3904 crosses the tile boundary at 3840, y = 2130..2190 the one at 2176; the second crop
    # sits on the silhouette of the refractive mesh
    for crop in ((3776, 2130, 3904, 2190), (3300, 1500, 3364, 1564)):
        ref = ref_scene.render(cam, q, crop=crop)["rgba8"]
        got = frame[crop[1]:crop[3], crop[0]:crop[2]].cpu().numpy()
        assert np.array_equal(got, ref), crop
    sc.close()


def test_c5b_256_instances_full_size(gpu_rt, ob):
    """Config C5b at its real size: one 327 680-triangle BLAS, 256 instances (top-level tree), 3840x2160, depth 8.
    Crops of the CUDA frame are bit-identical to the oracle."""
    rt = gpu_rt
    inputs = scenes.c5b_scene(rt, n=16, level=7, sky_size=512)
    cam = scenes.c5_camera(rt, 16)
    p = rt.FrameParams.make(3840, 2160, 8, 4)
    sc = rt.Scene(inputs, device_id=0)
    img = sc.render(cam, p)
    st = sc.stats()
    assert st["stack_overflow"] == 0 and st["rays_primary"] == 3840 * 2160 * 4
    ref_scene = ob.OracleScene(inputs)
    for crop in ((1888, 1060, 1952, 1100), (600, 1500, 664, 1532)):
        ref = ref_scene.render(cam, p, crop=crop)["rgba8"]
        assert np.array_equal(img[crop[1]:crop[3], crop[0]:crop[2]], ref), crop
    sc.close()


def test_c5a_84m_triangles_full_size(gpu_rt, ob):
    """Config C5a at its real size: 16 x 16 stand-in meshes flattened into ONE mesh of 83 886 080 triangles (the
    deepest tree and the longest traversal stacks of all configs), 3840x2160, depth 8.  The oracle cannot build a
    tree over 84 M triangles in test time, so the check is the independent float64 renderer on the SAME geometry
    expressed as 256 instances of the stand-in (vertex + translation; the flattened mesh holds that sum rounded to
    FP32, so ids are compared through the copy index and colours at BASELINE.json's thresholds), plus the
    bit-exact oracle on a 4 x 4 grid (5.2 M triangles) flattened the same way."""
    rt = gpu_rt
    n, level = 16, 7
    av, ai = rt.make_standin_armadillo(level)
    ntri = ai.shape[0]
    inputs = scenes.c5a_scene(rt, n=n, level=level, sky_size=512)
    cam = scenes.c5_camera(rt, n)
    p = rt.FrameParams.make(3840, 2160, 8, 4)
    sc = rt.Scene(inputs, device_id=0)
    img = sc.render(cam, p)
    st = sc.stats()
    assert st["stack_overflow"] == 0 and 1 < st["max_tree_depth"] <= 128
    prim = sc.primary_hits(cam, p).reshape(2160, 3840, 4)
    sc.close()
    del inputs
    # the same triangles as instances (copy k = iz * n + ix sits at the offset rth_make_grid adds)
    inst = [(0, rt.TYPE_DIFFUSE, rt.translation34((ix - 0.5 * (n - 1)) * scenes.GRID_PITCH, 0.0, (iz - 0.5 * (n - 1)) * scenes.GRID_PITCH))
            for iz in range(n) for ix in range(n)]
    inst_inputs = rt.SceneInputs([(av, ai)], inst, rt.make_procedural_skybox(512))
    o64 = ob.Oracle64Scene(inst_inputs)
    crop = (1700, 1000, 2100, 1200)
    ref = o64.render(cam, p, crop=crop)
    got = prim[crop[1]:crop[3], crop[0]:crop[2]].reshape(-1)
    hit = ref["primary"]["instance"] != 0xFFFFFFFF
    flat = np.where(hit, ref["primary"]["instance"].astype(np.int64) * ntri + ref["primary"]["primitive"], 0xFFFFFFFF)
    same = (got["primitive"].astype(np.int64) == flat) & ((got["instance"] == 0xFFFFFFFF) == ~hit)
    # (not the 99.99 % of identical geometry: the flattened mesh holds fl32(vertex + offset), the instances the exact
    # sum, and at this distance the triangles are smaller than a pixel, so a few 1e-4 of the samples sit within that
    # rounding of an edge; a lost subtree or a dropped stack entry would show up as whole missing triangles)
    assert hit.mean() > 0.3 and same.mean() >= 0.999, (hit.mean(), same.mean())
    d = np.abs(img[crop[1]:crop[3], crop[0]:crop[2]].astype(np.int32) - ref["rgba8"].astype(np.int32)).max(axis=2)
    assert (d > 2).mean() <= 1e-4 and psnr(img[crop[1]:crop[3], crop[0]:crop[2]], ref["rgba8"]) >= 45.0, ((d > 2).sum(), d.max())
    # bit-exact on a smaller flattened grid
    small = scenes.c5a_scene(rt, n=4, level=level, sky_size=64)
    cam4 = scenes.c5_camera(rt, 4)
    q = rt.FrameParams.make(960, 540, 8, 4)
    sc = rt.Scene(small, device_id=0)
    img4 = sc.render(cam4, q)
    sc.close()
    c4 = (400, 240, 560, 300)
    assert np.array_equal(img4[c4[1]:c4[3], c4[0]:c4[2]], ob.OracleScene(small).render(cam4, q, crop=c4)["rgba8"])


def assert_frame_against_oracle64(rt, ob, inputs, params, cam, crop, chaotic=False):
    sc = rt.Scene(inputs, device_id=0)
    img = sc.render(cam, params)
    prim = sc.primary_hits(cam, params).reshape(params.height, params.width, params.samples_per_pixel)
    sc.close()
    x0, y0, x1, y1 = crop
    s32 = ob.OracleScene(inputs)
    sig32 = ob.oracle32_signatures(s32, cam, params, crop=crop)   # the CUDA frame equals oracle32 bit for bit: its paths
    s64 = ob.Oracle64Scene(inputs)
    ref = s64.render(cam, params, crop=crop)
    rep = ob.compare_with_oracle64(img[y0:y1, x0:x1], prim[y0:y1, x0:x1].reshape(-1), ref, sig32)
    ob.explain_arithmetic_residuals(rep, s64, cam, params, crop, ref)
    assert rep["primary_match"] >= 0.9999 and rep["residual_primary"] == 0, rep
    if not chaotic:   # all of BASELINE.json's thresholds; the few pixels above 2/255 must each be explained
        assert rep["psnr"] >= 45.0 and rep["residual_unexplained"] == 0, rep
        assert rep["residual_pixels"] <= max(1, int(1e-4 * rep["pixels"])), rep
    else:   # refractive stand-in at depth 8: see tests/test_oracle64.py::test_c3_residuals_are_path_divergence
        assert rep["residual_pixels"] <= 0.015 * rep["pixels"], rep
        assert rep["residual_unexplained"] <= 2e-5 * rep["pixels"], rep
    return rep


def test_cuda_frames_against_the_independent_float64_renderer(gpu_rt, ob):
    """CUDA vs oracle64.cpp (written from the shaders without the FP32 contract) at BASELINE.json's thresholds:
    C1 whole frame, C2 at 1920x1080, a crop of C3 at 3840x2160 (residual pixels listed by cause in the report)."""
    rt = gpu_rt
    cam = rt.CameraVectors.default()
    r1 = assert_frame_against_oracle64(rt, ob, scenes.c1_scene(rt), scenes.c1_params(rt), cam, (0, 0, 800, 600))
    r2 = assert_frame_against_oracle64(rt, ob, scenes.c2_scene(rt, 2048), rt.FrameParams.make(1920, 1080, 4, 4), cam, (0, 0, 1920, 1080))
    r3 = assert_frame_against_oracle64(rt, ob, scenes.c3_scene(rt, level=7, sky_size=512), rt.FrameParams.make(3840, 2160, 8, 4), cam,
                                       (1536, 864, 2304, 1296), chaotic=True)
    print("oracle64 reports:", r1, r2, r3)


def deep_tree_mesh(rt, per_level=40, levels=21):
    """A mesh whose Morton-code tree is a long chain: clusters of nearly coincident triangles at the geometrically
    shrinking positions 8 * 2^-k along the diagonal, approaching the corner of the centroid bounds: every level of
    the 63-bit code peels one cluster off, and inside a cluster the low bits / the index tie-break add
    log2(per_level) more levels."""
    verts, idx = [], []
    for k in range(levels + 1):
        c = 8.0 * 2.0 ** -k if k < levels else 0.0
        s = 0.3 * 2.0 ** -min(k, levels - 1)
        for j in range(per_level):
            b = len(verts)
            e = s * (1.0 + 0.001 * j)   # slightly different shapes: every triangle is its own primitive
            for (dx, dy) in ((-e, -e), (2 * e, -e), (-e, 2 * e)):
                verts.append([c + dx, c + dy, c, 0.0, 0.0, 1.0])
            idx.append([b, b + 1, b + 2])
    return np.array(verts, np.float32), np.array(idx, np.uint32)


def test_deep_tree_is_traversed_completely(gpu_rt, ob):
    """Trees far deeper than the 8 stack entries kept in shared memory, up to the edge of what round 1's 64-entry
    stack held (the treelet pass turns clusters of coincident triangles into chains: 63 levels for this mesh): the
    build reports the depth, the traversal finds exactly the oracle's brute-force hits, no stack entry is dropped."""
    rt = gpu_rt
    rng = np.random.default_rng(3)
    for per_level, flags, min_depth, n in ((40, rt.FLAG_NO_REFINE, 24, 20000), (1000, 0, 40, 8000), (4096, 0, 56, 3000)):
        v, i = deep_tree_mesh(rt, per_level)
        inputs = rt.SceneInputs([(v, i)], [(0, rt.TYPE_DIFFUSE, rt.IDENTITY34)], None)
        sc = rt.Scene(inputs, device_id=0, flags=flags)
        depth = sc.stats()["max_tree_depth"]
        assert min_depth <= depth <= 128, (per_level, depth)
        # rays through the cluster centres (they pierce all the coincident triangles of a cluster) and the corner
        k = rng.integers(0, 21, n)
        tgt = (8.0 * 2.0 ** -k)[:, None] * np.ones((n, 3)) + rng.normal(0, 1, (n, 3)) * (0.1 * 2.0 ** -k)[:, None]
        tgt[::7] = rng.normal(0, 1e-6, (len(tgt[::7]), 3))
        o = np.tile(np.array([0.3, -0.2, 30.0]), (n, 1)) + rng.normal(0, 2.0, (n, 3))
        d = tgt - o
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        got = sc.trace_rays(o.astype(np.float32), d.astype(np.float32))
        assert sc.stats()["stack_overflow"] == 0
        ref = ob.OracleScene(inputs).trace_rays(o.astype(np.float32), d.astype(np.float32), brute=True)
        assert (ref["instance"] != 0xFFFFFFFF).mean() > 0.5
        for f in ("instance", "primitive", "t", "u", "v"):
            assert np.array_equal(got[f], ref[f]), (per_level, f)
        any_got = sc.trace_rays(o.astype(np.float32), d.astype(np.float32), any_hit=True)
        assert np.array_equal(any_got["instance"] != 0xFFFFFFFF, ref["instance"] != 0xFFFFFFFF)
        sc.close()
        if per_level == 1000:   # and a frame through the wavefront pipeline (shadow rays included)
            check_frame(rt, ob, inputs, rt.FrameParams.make(96, 64, 2, 2), cam=scenes.look_at_camera(rt, (4.0, 3.0, 30.0), (1.0, 1.0, 1.0)))


def run_gpu_python(code, env):
    import subprocess, sys
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, "-c", code], env=e, capture_output=True, text=True, timeout=900,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


def test_too_deep_tree_is_refused_at_build_time(gpu_rt):
    """RT_DEPTH_LIMIT lowers the depth the traversal stacks are declared to hold: rt_scene_build must then fail
    closed with RT_ERR_UNSUPPORTED instead of rendering with dropped stack entries."""
    out = run_gpu_python("""
import sys; sys.path.insert(0, 'tests')
import __graft_entry__ as g
rt = g.load_package()
import test_gpu_parity as t
v, i = t.deep_tree_mesh(rt)
try:
    rt.Scene(rt.SceneInputs([(v, i)], [(0, rt.TYPE_DIFFUSE, rt.IDENTITY34)], None), device_id=0, flags=rt.FLAG_NO_REFINE)
    print('BUILT')
except rt.RtError as e:
    print('REFUSED', e.code, str(e))
""", {"RT_DEPTH_LIMIT": "12"})
    assert "REFUSED" in out and str(gpu_rt.ERR_UNSUPPORTED) in out and "levels deep" in out, out


PARITY_SNIPPET = """
import sys; sys.path.insert(0, 'tests')
import __graft_entry__ as g
rt = g.load_package()
import os, oracle_binding as ob, scenes, test_gpu_parity as t
A = int(os.environ.get('TEST_ACCEL', '0'))
t.check_frame(rt, ob, scenes.smoke_scene(rt), rt.FrameParams.make(160, 120, 6, 4), A)
t.check_frame(rt, ob, scenes.c3_scene(rt, level=5, sky_size=256), rt.FrameParams.make(384, 216, 8, 4), A)
t.check_frame(rt, ob, scenes.c1_like_scene(rt), rt.FrameParams.make(203, 151, 63, 4), A)
t.check_frame(rt, ob, scenes.instanced_grid_scene(rt), rt.FrameParams.make(240, 160, 5, 4), A)     # 13 instances: top-level tree
sc = rt.Scene(scenes.c3_scene(rt, level=6, sky_size=128), device_id=0, accel=A)
p = rt.FrameParams.make(2048, 1600, 8, 4)          # 13.1 M sample slots: above the two-sub-frame threshold
img = sc.render(rt.CameraVectors.default(), p)
crop = (960, 760, 1088, 824)
ref = ob.OracleScene(scenes.c3_scene(rt, level=6, sky_size=128)).render(rt.CameraVectors.default(), p, crop=crop)['rgba8']
assert (img[crop[1]:crop[3], crop[0]:crop[2]] == ref).all()
print('PARITY OK', sc.stats()['kernel_launches'])
"""


@pytest.mark.parametrize("env", [
    {"RT_FUSED": "1"},                                   # bounce-fused path kernel (rt_path.cuh)
    {"RT_FUSED": "1", "RT_FUSE_FROM": "2"},              # wavefront launches for bounces 0-1, fused kernel after
    {"RT_FUSED": "1", "RT_PATH_SERVE": "1", "RT_PATH_LEAF": "1", "RT_PATH_CHUNK": "32"},
    {"RT_FUSED": "1", "RT_PATH_SERVE": "32", "RT_PATH_LEAF": "32", "RT_SUBFRAMES": "1"},
    {"RT_MIGRATE": "4"},                                 # stragglers of a batch set aside with their traversal state, resumed together
    {"RT_MIGRATE": "12", "RT_SUBFRAMES": "1"},
    {"RT_MIGRATE": "1", "RT_SUBFRAMES": "3", "RT_SUBFRAMES_MIN": "1"},
    {"RT_TLAS_STEP": "0", "RT_SHADOW_TLAS_STEP": "0"},   # top-level tree walked per lane (trace_scene) instead of with the warp in step
    {"RT_TAIL_FROM": "0"},                               # deep bounce limits bounce by bounce instead of the path-kernel tail
    {"RT_TAIL_FROM": "3"},
    {"RT_SHADOW_STEP": "1"},                             # shadow rays through looped instances with the warp in step
    {"RT_INSTANCE_STEP": "0"},                           # the generic closest-hit + shading launch (trace_scene) for looped instances
    {"RT_FUSE_SHADE": "0"},                              # separate shading kernels between the traversal launches (round 1)
    {"RT_FUSE_SHADE": "0", "RT_SHADOW_STREAM": "0", "RT_SUBFRAMES": "1"},
    {"RT_FUSE_SHADE": "0", "TEST_ACCEL": "4"},
    {"RT_SUBFRAMES": "1"},
    {"RT_SUBFRAMES": "3", "RT_SUBFRAMES_MIN": "1"},
    {"RT_SHADOW_STREAM": "0"},
    {"RT_EARLY_RESOLVE": "0"},
    {"RT_SORT_FROM": "1", "RT_SORT_TO": "8"},            # ray reordering between bounces
    {"RT_WAVE": "1", "TEST_ACCEL": "3"},                 # warp-synchronous traversal of the 8-wide layout (rt_wave.cuh)
    {"TEST_ACCEL": "3", "RT_SUBFRAMES": "1"},
    {"TEST_ACCEL": "4"},                                 # 4-wide full-precision nodes (RT_ACCEL_WIDE4)
], ids=lambda e: ",".join(f"{k}={v}" for k, v in e.items()))
def test_switchable_paths_are_parity_green(gpu_rt, env):
    """Every run-time switch of the frame pipeline renders the same bits as the default path:
    the switches are read once per process, so each setting runs in its own interpreter."""
    assert "PARITY OK" in run_gpu_python(PARITY_SNIPPET, env)


def test_repeated_large_frames_are_bit_identical(gpu_rt):
    """compute-sanitizer is closed on this pool (profiles/r02_sanitizer.txt), so the race check of the multi-stream
    frame (two sub-frames, shadow rays on their own streams, high-priority early resolve, cross-stream events) is
    empirical: a frame above the two-sub-frame threshold rendered ten times, with device work of another stream
    interleaved, must give the same bits every time - every sample slot is written by exactly one path, so any
    ordering bug between the streams shows up as a changed pixel."""
    import hashlib, torch
    rt = gpu_rt
    sc = rt.Scene(scenes.c3_scene(rt, level=6, sky_size=128), device_id=0)
    cam = rt.CameraVectors.default()
    p = rt.FrameParams.make(2048, 1600, 8, 4, output_location=rt.LOCATION_DEVICE)
    out = torch.zeros((1600, 2048, 4), dtype=torch.uint8, device="cuda:0")
    noise = torch.zeros(64 << 20, dtype=torch.uint8, device="cuda:0")
    side = torch.cuda.Stream()
    hashes = set()
    for i in range(10):
        out.zero_()
        torch.cuda.synchronize()
        with torch.cuda.stream(side):      # unrelated work competing for SMs and copy engines, different each time
            for _ in range(i % 4):
                noise.add_(1)
        sc.render_into(cam, p, out.data_ptr())
        torch.cuda.synchronize()
        hashes.add(hashlib.sha256(out.cpu().numpy().tobytes()).hexdigest())
        assert sc.stats()["stack_overflow"] == 0
    assert len(hashes) == 1
    sc.close()


def test_sky_texture_object_within_tolerance(gpu_rt, ob):
    """RT_FLAG_SKY_TEXTURE (north_star (3); the VkSampler of main.cpp:2376-2412 as a CUDA texture object): the hardware
    filter uses 8-bit weights, so the frame may differ from the exact path - by at most 2/255 (BASELINE.json), PSNR >= 45 dB,
    and the traced rays are the same."""
    rt = gpu_rt
    cam = rt.CameraVectors.default()
    for inputs, p in ((scenes.c2_scene(rt, 256), rt.FrameParams.make(480, 270, 4, 4)),
                      (scenes.c1_scene(rt), rt.FrameParams.make(400, 300, 63, 4)),
                      (scenes.smoke_scene(rt), rt.FrameParams.make(160, 120, 6, 4))):
        exact = rt.Scene(inputs, device_id=0)
        a = exact.render(cam, p)
        sa = exact.stats()
        exact.close()
        tex = rt.Scene(inputs, device_id=0, flags=rt.FLAG_SKY_TEXTURE)
        b = tex.render(cam, p)
        sb = tex.stats()
        tex.close()
        d = np.abs(a.astype(np.int32) - b.astype(np.int32))
        assert d.max() <= 2 and psnr(a, b) >= 45.0, (d.max(), psnr(a, b))
        assert np.array_equal(a, ob.OracleScene(inputs).render(cam, p)["rgba8"])
        for k in ("rays_primary", "rays_secondary", "rays_shadow", "n_miss", "n_hit"):
            assert sa[k] == sb[k]
