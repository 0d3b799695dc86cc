"""N > 1 path on CPU: two processes over gloo shard a frame by interleaved tiles exactly as the
C ABI defines it (rt_frame_params.shard_index/shard_count), gather the per-shard buffers on rank 0
and de-interleave them like rt_assemble_shards; the result must equal the unsharded oracle
frame bit for bit.  Pixels are produced by the CPU emulation of the device functions (tests
only), so this checks the tile ownership / layout / gather logic, not the GPU."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def shard_pixels(w, h, tw, th, n):
    tiles = ((w + tw - 1) // tw) * ((h + th - 1) // th)
    return ((tiles + n - 1) // n) * tw * th


def assemble(gathered, w, h, tw, th, n):
    """numpy mirror of k_assemble (rt_render.cu)."""
    tiles_x = (w + tw - 1) // tw
    y, x = np.mgrid[0:h, 0:w]
    tile = (y // th) * tiles_x + x // tw
    src = (tile // n) * (tw * th) + (y % th) * tw + (x % tw)
    return gathered[tile % n, src]


def worker(rank, world, port, w, h, tw, th, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import __graft_entry__ as graft
    rt = graft.load_package()
    import emu_binding as emu
    import scenes
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    inputs = scenes.smoke_scene(rt)
    cam = rt.CameraVectors.default()
    p = rt.FrameParams.make(w, h, 4, 2, tile_w=tw, tile_h=th, shard_index=rank, shard_count=world)
    full = emu.EmuScene(inputs, 0).render(cam, p)["rgba8"]  # global coordinates, global launch size
    per = shard_pixels(w, h, tw, th, world)
    local = np.zeros((per, 4), np.uint8)
    tiles_x = (w + tw - 1) // tw
    tiles = tiles_x * ((h + th - 1) // th)
    n_local = (tiles - rank + world - 1) // world if tiles > rank else 0
    for slot in range(n_local * tw * th):
        ok, x, y = emu.slot_to_pixel(p, slot)
        if ok:  # compact shard layout: tile after tile, each tile row-major (rt_b200.h)
            tile = (y // th) * tiles_x + x // tw
            local[(tile // world) * tw * th + (y % th) * tw + (x % tw)] = full[y, x]
    t = torch.from_numpy(local)
    gathered = [torch.zeros_like(t) for _ in range(world)] if rank == 0 else None
    dist.gather(t, gathered, dst=0)
    if rank == 0:
        frame = assemble(np.stack([g.numpy() for g in gathered]), w, h, tw, th, world)
        np.save(out_path, frame)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("w,h,tw,th", [(100, 68, 32, 16), (64, 48, 8, 4)])
def test_two_rank_tile_sharding_matches_full_frame(rt, ob, tmp_path, w, h, tw, th):
    import scenes
    out = str(tmp_path / "frame.npy")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(worker, args=(2, port, w, h, tw, th, out), nprocs=2, join=True)
    frame = np.load(out)
    ref = ob.OracleScene(scenes.smoke_scene(rt)).render(rt.CameraVectors.default(), rt.FrameParams.make(w, h, 4, 2))["rgba8"]
    assert np.array_equal(frame, ref)


def in_frame_worker(rank, world, port, w, h, tw, th, frame_path):
    """RT_SHARD_IN_FRAME on CPU: every rank stores its tiles straight into one shared frame (a memmap stands in for
    the peer-mapped device frame), in the row-by-row order of k_resolve, then a barrier; no gather."""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import __graft_entry__ as graft
    rt = graft.load_package()
    import emu_binding as emu
    import scenes
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    inputs = scenes.smoke_scene(rt)
    cam = rt.CameraVectors.default()
    p = rt.FrameParams.make(w, h, 4, 2, tile_w=tw, tile_h=th, shard_index=rank, shard_count=world)
    full = emu.EmuScene(inputs, 0).render(cam, p)["rgba8"]
    frame = np.memmap(frame_path, dtype=np.uint8, mode="r+", shape=(h, w, 4))
    tiles_x = (w + tw - 1) // tw
    tiles = tiles_x * ((h + th - 1) // th)
    n_local = (tiles - rank + world - 1) // world if tiles > rank else 0
    for q in range(n_local * tw * th):          # q walks each tile row-major, like the in-frame branch of k_resolve
        lt, inn = divmod(q, tw * th)
        ty, tx = divmod(inn, tw)
        pslot = lt * tw * th + ((ty // 4) * (tw // 8) + tx // 8) * 32 + (ty % 4) * 8 + tx % 8
        ok, x, y = emu.slot_to_pixel(p, pslot)  # the slot whose samples belong to that pixel
        tile = lt * world + rank
        assert (x, y) == ((tile % tiles_x) * tw + tx, (tile // tiles_x) * th + ty)
        if ok:
            frame[y, x] = full[y, x]
    frame.flush()
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_store_into_one_frame(rt, ob, tmp_path):
    import scenes
    w, h, tw, th = 100, 68, 32, 16
    path = str(tmp_path / "frame.bin")
    np.full((h, w, 4), 7, np.uint8).tofile(path)
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(in_frame_worker, args=(2, port, w, h, tw, th, path), nprocs=2, join=True)
    frame = np.fromfile(path, np.uint8).reshape(h, w, 4)
    ref = ob.OracleScene(scenes.smoke_scene(rt)).render(rt.CameraVectors.default(), rt.FrameParams.make(w, h, 4, 2))["rgba8"]
    assert np.array_equal(frame, ref)


def host_frame_worker(rank, world, port, w, h, tw, th, out_path):
    """rt_host_frame_* protocol on CPU: the frame and its arrival flags live in POSIX shared memory (what
    rt_host_frame_create maps and registers with CUDA on a GPU box), every rank stores its tiles in place and
    reports the frame number with rt_host_frame_signal; rank 0 reads the frame as soon as rt_host_frame_wait
    returns - the flags are the only synchronisation between the stores and the read.  Three frames with
    different contents go through the same memory."""
    import ctypes as C
    import gc
    from multiprocessing import shared_memory
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import __graft_entry__ as graft
    rt = graft.load_package()
    import emu_binding as emu
    import scenes
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nbytes = w * h * 4
    total = ((nbytes + 255) & ~255) + 4096            # frame, then the flags (rt_api.cu: peer_flag_offset)
    names = [None]
    shm = None
    if rank == 0:
        shm = shared_memory.SharedMemory(create=True, size=total)
        shm.buf[:total] = bytes(total)
        names = [shm.name]
    dist.broadcast_object_list(names, src=0)
    if rank != 0:
        shm = shared_memory.SharedMemory(name=names[0])
    anchor = C.c_char.from_buffer(shm.buf)
    hf = rt.HostFrame(names[0], 0, C.addressof(anchor), 0, nbytes, False)
    frame = np.frombuffer(shm.buf, np.uint8, nbytes).reshape(h, w, 4)
    inputs = scenes.smoke_scene(rt)
    cam = rt.CameraVectors.default()
    p = rt.FrameParams.make(w, h, 4, 2, tile_w=tw, tile_h=th, shard_index=rank, shard_count=world)
    full = emu.EmuScene(inputs, 0).render(cam, p)["rgba8"]
    tiles_x = (w + tw - 1) // tw
    tiles = tiles_x * ((h + th - 1) // th)
    n_local = (tiles - rank + world - 1) // world if tiles > rank else 0
    mine = []
    for q in range(n_local * tw * th):
        lt, inn = divmod(q, tw * th)
        ty, tx = divmod(inn, tw)
        tile = lt * world + rank
        x, y = (tile % tiles_x) * tw + tx, (tile // tiles_x) * th + ty
        if x < w and y < h:
            mine.append((y, x))
    ys, xs = np.array(mine).T
    got = []
    for k in (1, 2, 3):
        dist.barrier()                                 # the consumer is done with frame k-1 (bench.py does the same)
        frame[ys, xs] = full[ys, xs] + np.uint8(k)     # this rank's tiles of frame k
        hf.signal(rank, k)
        if rank == 0:
            hf.wait(world, k, timeout_ms=20000)
            got.append(frame.copy())
    if rank == 0:
        np.save(out_path, np.stack(got))
    dist.barrier()
    hf.host_ptr = 0
    del frame, anchor, hf
    gc.collect()
    shm.close()
    if rank == 0:
        shm.unlink()
    dist.destroy_process_group()


def test_two_ranks_deliver_frames_through_shared_host_memory_and_flags(rt, ob, tmp_path):
    import scenes
    w, h, tw, th = 100, 68, 32, 16
    out = str(tmp_path / "frames.npy")
    port = 33500 + (os.getpid() % 2000)
    mp.spawn(host_frame_worker, args=(2, port, w, h, tw, th, out), nprocs=2, join=True)
    frames = np.load(out)
    ref = ob.OracleScene(scenes.smoke_scene(rt)).render(rt.CameraVectors.default(), rt.FrameParams.make(w, h, 4, 2))["rgba8"]
    for k in (1, 2, 3):
        assert np.array_equal(frames[k - 1], ref + np.uint8(k))
