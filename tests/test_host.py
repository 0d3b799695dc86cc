"""Host-side logic (no GPU): asset packing against the reference's own loader output, Camera
against the reference's unmodified camera.cpp, the C-ABI libraries load and export every symbol
their headers declare, and the product refuses to run without a CUDA device."""
import ctypes as C
import hashlib
import os
import re

import numpy as np
import pytest

import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def declared_functions(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rth?_[a-z0-9_]+)\s*\(", src)))


def test_rt_b200_exports_every_declared_symbol(rt):
    lib = C.CDLL(rt.RT_LIB_PATH)
    names = declared_functions("rt_b200.h")
    assert len(names) >= 17
    for n in names:
        assert hasattr(lib, n), f"librt_b200.so does not export {n}"
    lib.rt_abi_version.restype = C.c_uint32
    assert lib.rt_abi_version() == 2
    assert lib.rt_device_count() >= 0


def test_rt_host_exports_every_declared_symbol(rt):
    lib = C.CDLL(rt.HOST_LIB_PATH)
    names = declared_functions("rt_host.h")
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"librt_host.so does not export {n}"


def test_struct_layouts_match_header(rt):
    assert C.sizeof(rt.MeshDesc) == 32 and C.sizeof(rt.InstanceDesc) == 56
    assert C.sizeof(rt.CameraVectors) == 48 and C.sizeof(rt.FrameParams) == 60 and C.sizeof(rt.Hit) == 20 and C.sizeof(rt.IpcHandle) == 64
    assert rt.HIT_DTYPE.itemsize == 20
    assert C.sizeof(rt.Stats) == 152 and C.sizeof(rt.Material) == 44


def test_no_device_means_error_not_fallback(rt):
    if rt.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(rt.RtError, match="no CUDA device"):
        rt.Scene(scenes.smoke_scene(rt))
    with pytest.raises(rt.RtError):
        rt.sort_pairs(np.arange(4, dtype=np.uint64), np.arange(4, dtype=np.uint32))


def test_host_frame_without_a_device_and_its_arrival_flags(rt):
    """rt_host_frame_create needs a CUDA device (the frame is registered with it) and leaves no shared-memory
    name behind when there is none; the arrival flags behind the frame are plain host code: ranks report a
    frame number, the wait returns once all have reached it (older numbers count as reached) and fails after
    its timeout when one is missing."""
    name = f"/rt_b200_cpu_test_{os.getpid()}"
    if rt.device_count() == 0:
        with pytest.raises(rt.RtError, match="no CUDA device"):
            rt.HostFrame.create(name, 0, 4096)
        assert not os.path.exists("/dev/shm" + name)
    nbytes = 1000                                         # the flags start at the next 256-byte boundary
    buf = np.zeros(1024 + 4096, np.uint8)
    hf = rt.HostFrame(name, 0, buf.ctypes.data, 0, nbytes, False)
    for r in (2, 0, 1):
        hf.signal(r, 7)
    hf.wait(3, 7)
    hf.wait(3, 6)
    with pytest.raises(rt.RtError, match="did not deliver"):
        hf.wait(4, 7, timeout_ms=20)
    flags = buf[1024:1024 + 16].view(np.uint32)
    assert flags.tolist() == [7, 7, 7, 0] and not buf[:1024].any()
    with pytest.raises(rt.RtError):
        hf.signal(5000, 1)                                # more ranks than flags
    hf.host_ptr = 0                                       # nothing to unmap: the buffer is numpy's


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout absent")
@pytest.mark.parametrize("name", ["cube", "teapot"])
def test_loader_reproduces_reference_packing(rt, name):
    """rth_load_obj vs the golden buffers packed by the reference's tiny_obj_loader.h + the
    rules of main.cpp:1638-1700 (tests/golden/make_golden.py)."""
    v, i, policy = rt.load_obj(f"{REF}/resources/{name}.obj")
    gv, gi = scenes.golden_mesh(name)
    assert policy == 0
    assert np.array_equal(v, gv) and np.array_equal(i, gi)


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout absent")
def test_malformed_cube_scene_uses_corrected_normals(rt):
    v, i, policy = rt.load_obj(f"{REF}/resources/cube_scene.obj")
    assert policy == 1 and i.shape == (42, 3) and i.max() < v.shape[0]
    assert np.allclose(np.linalg.norm(v[:, 3:], axis=1), 1.0, atol=1e-3)


def test_obj_parser_variants(rt, tmp_path):
    p = tmp_path / "quad.obj"
    p.write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nf 1 2 3 4\nf -4 -3 -2\n")
    v, i, policy = rt.load_obj(str(p))
    assert policy == 2 and i.tolist() == [[0, 1, 2], [0, 2, 3], [0, 1, 2]]
    assert np.allclose(v[:, 3:], [0, 0, 1])
    with pytest.raises(rt.RtError):
        rt.load_obj(str(tmp_path / "missing.obj"))


def test_camera_matches_reference_camera(rt):
    """Scripted input sequence replayed on the product's Camera; golden values come from the
    reference's own camera.cpp compiled with a GLM shim (oracle/ref_build.sh)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(scenes.GOLDEN, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    gold = np.load(os.path.join(scenes.GOLDEN, "camera_golden.npz"))
    cam = rt.Camera((0.0, 0.0, 20.0))
    snaps = [np.concatenate(cam._get())]
    views = [np.concatenate([cam.getViewingMatrix().T.ravel(), cam.getViewingMatrix(True).T.ravel()])]
    for op in mg.CAMERA_SCRIPT:
        if op[0] == "mouse":
            cam.processMouseMovement(op[1], op[2])
        elif op[0] == "move":
            cam.move(op[1], op[2])
        else:
            cam.look(op[1])
        snaps.append(np.concatenate(cam._get()))
        views.append(np.concatenate([cam.getViewingMatrix().T.ravel(), cam.getViewingMatrix(True).T.ravel()]))
    assert np.array_equal(np.array(snaps, np.float32), gold["vectors"])
    assert np.allclose(np.array(views, np.float32), gold["views"], atol=1e-6)
    # default pose: position (0,0,20), looking down -Z (camera.cpp:8-25)
    assert np.allclose(gold["vectors"][0], [0, 0, 20, 1, 0, 0, 0, 1, 0, 0, 0, -1], atol=1e-7)


def test_orbit_transforms(rt):
    x0 = rt.orbit_transforms(0.0)
    assert np.allclose(x0[0], rt.IDENTITY34) and np.allclose(x0[1], rt.translation34(0, 0, 5), atol=1e-6)
    x = rt.orbit_transforms(1.0)  # half a turn: T(0,0,-5) R_y(pi) T(0,0,10) = rotation + T(0,0,-15)
    assert np.allclose(x[1].reshape(3, 4)[:, 3], [0, 0, -15], atol=1e-5)


def test_standin_mesh_and_grid_are_deterministic(rt):
    v, i = rt.make_standin_armadillo(3)
    assert v.shape == (642, 6) and i.shape == (1280, 3)
    v2, i2 = rt.make_standin_armadillo(3)
    assert np.array_equal(v, v2) and np.array_equal(i, i2)
    assert np.allclose(np.linalg.norm(v[:, 3:], axis=1), 1.0, atol=1e-5)
    r = np.linalg.norm(v[:, :3], axis=1)
    assert 2.5 * 0.84 < r.min() and r.max() < 2.5 * 1.16
    gv, gi = rt.make_grid(v, i, 3, 2, 7.0)
    assert gv.shape == (642 * 6, 6) and gi.shape == (1280 * 6, 3) and gi.max() == 642 * 6 - 1
    assert np.allclose(gv[:642, 0] + 7.0, gv[642:1284, 0])


def test_procedural_skybox_and_ppm(rt, tmp_path):
    sky = rt.make_procedural_skybox(32)
    assert sky.shape == (6, 32, 32, 4) and (sky[..., 3] == 255).all()
    assert np.array_equal(sky, rt.make_procedural_skybox(32))
    img = sky[4]
    rt.write_ppm(str(tmp_path / "a.ppm"), img)
    raw = open(tmp_path / "a.ppm", "rb").read()
    assert raw.startswith(b"P6\n32 32\n255\n") and raw[13:] == img[..., :3].tobytes()


def test_frame_shard_sizes(rt):
    lib = rt.rt_lib()
    p = rt.FrameParams.make(200, 136, 1, 1, tile_w=32, tile_h=16, shard_count=3)
    tiles = 7 * 9
    assert lib.rt_frame_shard_pixels(C.byref(p)) == ((tiles + 2) // 3) * 512
    assert lib.rt_frame_output_bytes(C.byref(p)) == ((tiles + 2) // 3) * 512 * 4
    q = rt.FrameParams.make(200, 136, 1, 1, output_format=rt.FORMAT_RGBA32F)
    assert lib.rt_frame_output_bytes(C.byref(q)) == 200 * 136 * 16


HEADLESS = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "vulkan-raytracing_b200", "rt_headless")


def test_headless_host_usage_and_no_device(rt):
    """rt_headless (the C++ replacement of the reference's main()) is built, documents the config.h
    options, and refuses to run without a CUDA device instead of falling back."""
    import subprocess
    r = subprocess.run([HEADLESS, "--help"], capture_output=True, text=True)
    assert r.returncode == 0
    for opt in ("--center", "--orbiting", "--center-type", "--skybox", "--bounces", "--spp", "--fps"):
        assert opt in r.stdout
    assert subprocess.run([HEADLESS, "--bogus"], capture_output=True).returncode == 1
    if rt.device_count() == 0:
        r = subprocess.run([HEADLESS, "--center", "standin:2", "--orbiting", "standin:2", "--skybox", "procedural:8"],
                           capture_output=True, text=True)
        assert r.returncode == 2 and "no CPU fallback" in r.stderr


def test_mtl_reader(rt, tmp_path):
    """rth_load_mtl: every newmtl block becomes a record; absent statements keep the shader constants."""
    p = tmp_path / "m.mtl"
    p.write_text("# comment\nnewmtl red\nNs 17.5\nKa 0.3 0.1 0.1\nKd 0.9 0.3 0.2\nKs 0.5 0.5 0.9\nNi 1.33\nd 1.0\nillum 2\n"
                 "\nnewmtl bare\nKe 0 0 0\n")
    mats = rt.load_mtl(str(p))
    assert [n for n, _ in mats] == ["red", "bare"]
    red, bare = mats[0][1], mats[1][1]
    assert np.allclose(list(red.ka), [0.3, 0.1, 0.1]) and np.allclose(list(red.kd), [0.9, 0.3, 0.2])
    assert np.allclose(list(red.ks), [0.5, 0.5, 0.9]) and red.ns == np.float32(17.5) and red.ni == np.float32(1.33)
    d = rt.Material.make()
    assert bytes(bare) == bytes(d)
    with pytest.raises(RuntimeError):
        rt.load_mtl(str(tmp_path / "missing.mtl"))


def test_reference_mtl_files_parse(rt):
    """the .mtl files the reference ships (ignored by its shaders): Blender exports with Ns 225 / Ni 1.45"""
    root = "/root/reference/resources"
    if not os.path.isdir(root):
        pytest.skip("reference tree not present")
    mats = rt.load_mtl(os.path.join(root, "teapot.mtl"))
    assert len(mats) == 1 and mats[0][0] == "teapot" and mats[0][1].ns == 225.0 and mats[0][1].ni == np.float32(1.45)
    assert len(rt.load_mtl(os.path.join(root, "cube_scene.mtl"))) == 8


def test_host_normals_equal_a_numpy_restatement(rt):
    """rth_generate_normals: area-weighted, summed per vertex in triangle order in double (np.add.at keeps
    that order), normalised in double, rounded once; unused vertices get (0,1,0)."""
    v, i = rt.make_standin_armadillo(3)
    v = np.vstack([v, np.array([[9, 9, 9, 0, 0, 0]], np.float32)])       # a vertex no triangle uses
    i = np.vstack([i, np.array([[0, 0, 1]], np.uint32)])                  # a degenerate triangle
    got = rt.generate_normals_host(v, i)
    pos = v[:, :3].astype(np.float64)
    a, b, c = pos[i[:, 0]], pos[i[:, 1]], pos[i[:, 2]]
    n = np.cross(b - a, c - a)
    acc = np.zeros_like(pos)
    np.add.at(acc, i.reshape(-1), np.repeat(n, 3, axis=0))
    l = np.sqrt(acc[:, 0] * acc[:, 0] + acc[:, 1] * acc[:, 1] + acc[:, 2] * acc[:, 2])
    ref = np.where(l[:, None] == 0, np.array([0.0, 1.0, 0.0]), acc / np.where(l == 0, 1, l)[:, None]).astype(np.float32)
    assert np.array_equal(got[:, 3:], ref)
    assert np.array_equal(got[:, :3], v[:, :3])
    assert np.array_equal(got[-1, 3:], [0, 1, 0])


def test_committed_ncu_capture_belongs_to_the_committed_kernels():
    """profiles/traffic.json (DRAM bytes, warp / thread instructions of one C3 frame's traversal launches, from ncu) carries
    the hash of the CUDA sources it was captured from; bench.py prints roofline.traffic / issue_frac / simd_lanes only when
    that hash is the hash of the sources it runs.  A change of a kernel without a new capture shows up here."""
    import json
    import sys
    sys.path.insert(0, ROOT)
    import bench
    t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    assert t["source_hash"] == bench.source_hash()
    assert t["launches"] == 36 and t["k_trace_dram_bytes_per_frame"] > 0
