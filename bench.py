#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 ray-tracing core (contract in the task statement).

A "step" is one frame of the hot path (rt_render_frame) on a synthetic scene:
  N = 1   BASELINE config C3: refractive armadillo (stand-in mesh, 327 680 triangles) in the
          orbiting slot + diffuse teapot in the centre slot, 3840x2160, 4 spp, depth 8.
  N > 1   BASELINE config C4: the same scene at 7680x4320, tile-sharded over the N GPUs (scene
          replicated, one process per GPU), per-GPU tiles gathered over NCCL and assembled on
          rank 0.
metric: Mrays/s = traceRayEXT-equivalents (closest-hit + shadow traces) issued per second,
whole job.  `value` is timed with the frame left in HBM; `e2e` is the same metric through the
C ABI with a HOST output buffer (uniforms H2D + frame D2H inside the timed region).

`--impl reference` times the reference's CPU implementation of the path: the reference has no
CPU ray tracer and cannot be built here, so this is the oracle port (oracle/liboracle.so) on
all host cores, each step a bounded sample of the same frame: every k-th 64x64 tile (the tiles
of shard 0 of k in the C ABI's own tile sharding), i.e. an evenly spread 1/k of the frame with
the frame's own mix of sky, diffuse and refractive pixels.  The GPU renders exactly that shard
too (`same_rays`), so the CPU and GPU rates in one line are over the same rays.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# kernels are loaded when the library is, not at their first launch: `build_ms` times the build, not the driver's lazy loader
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")
ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import __graft_entry__ as graft  # noqa: E402

WORKLOADS = {
    # name: (scene builder name, width, height, depth, spp)
    "c3": ("c3_scene", 3840, 2160, 8, 4),
    "c4": ("c3_scene", 7680, 4320, 8, 4),
    "c2": ("c2_scene", 1920, 1080, 4, 4),
    "c5a": ("c5a_scene", 3840, 2160, 8, 4),
    "c5b": ("c5b_scene", 3840, 2160, 8, 4),
    "c1": ("c1_bench_scene", 800, 600, 63, 4),
}
WORKLOAD_TEXT = {
    "c3": "C3 armadillo_standin(327680 tri, refractive, T(0,0,5)) + teapot(diffuse, shadow rays) 3840x2160 spp4 depth8",
    "c4": "C4 armadillo_standin(327680 tri, refractive) + teapot(diffuse) 7680x4320 spp4 depth8, tile-sharded",
    "c2": "C2 teapot(mirror) + teapot(refractive, T(0,0,5)) 1920x1080 spp4 depth4",
    "c5a": "C5a 16x16 armadillo_standin grid flattened into one mesh (83.9 M tri, diffuse) 3840x2160 spp4 depth8, camera pulled back",
    "c5b": "C5b one armadillo_standin BLAS x 256 instances (diffuse/refractive alternating) 3840x2160 spp4 depth8, camera pulled back",
    "c1": "C1 cube_scene.obj(mirror, centre) + cube.obj(diffuse, T(0,0,5)), skybox_texture_sea fixture 256x256, 800x600 spp4 MAX_BOUNCE_COUNT 63 (config.h defaults)",
}
CPU_SAMPLE_RAYS = 5.5e6  # the CPU arms render every k-th tile, k chosen so that a step is about this many rays


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--accel", type=int, default=0, help="RT_ACCEL_*: 0 = default (binary BVH), 3 = compressed 8-wide BVH")
    ap.add_argument("--flags", type=int, default=0, help="extra rt_scene_desc flags, e.g. 16 = RT_FLAG_NO_REFINE")
    ap.add_argument("--sky", type=int, default=2048, help="procedural skybox face size")
    ap.add_argument("--e2e", default="host-frame", choices=["host-frame", "copy"],
                    help="how the frame reaches host memory in the e2e region: host-frame = every rank's k_resolve stores its "
                         "tiles into one shared pinned host frame (rt_host_frame_*), copy = frame gathered on rank 0's GPU, "
                         "then one device->host copy")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="N > 1: 'peer' = every rank's resolve kernel stores its tiles into rank 0's frame over NVLink "
                         "(falls back to 'nccl' if the frame cannot be shared); 'nccl' = gather collective + de-interleave")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, str(gpu_index)

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", self.gpu], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for k, n in enumerate(names):
                    if r[5 + k].lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def sample_params(rt, params, rays_per_frame):
    """rt_frame_params of the CPU sample: shard 0 of k of the frame's 64x64 tiles."""
    k = max(2, int(round(rays_per_frame / CPU_SAMPLE_RAYS)))
    q = rt.FrameParams.make(params.width, params.height, params.max_bounce_count, params.samples_per_pixel,
                            shard_index=0, shard_count=k)
    return q, k


def cpu_arm(rt, inputs, cam, params, steps, warmup, rays_per_frame, want_image=False):
    """Oracle port on all host cores over every k-th tile; returns (Mrays/s, ms/step, cores, sample text, rays, image)."""
    import oracle_binding as ob
    osc = ob.OracleScene(inputs)
    q, k = sample_params(rt, params, rays_per_frame)
    cores = ob.lib().orc_num_threads()
    rays, times, img = 0, [], None
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        img, c = osc.render_shard(cam, q, want_image=want_image)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
            rays = c["primary"] + c["secondary"] + c["shadow"]
    total = sum(times)
    sample = (f"every {k}th 64x64 tile of the {params.width}x{params.height} frame (shard 0 of {k} of the C ABI's tile sharding: "
              f"an evenly spread 1/{k} of the frame with its own ray mix), {rays} rays per step, oracle BVH mode, OpenMP over tile rows")
    return rays * len(times) / total / 1e6, 1e3 * total / len(times), cores, sample, rays, img


RAYS_PER_FRAME_HINT = {"c1": 3.0e6, "c3": 49.1e6, "c4": 196.3e6, "c2": 10.4e6, "c5a": 41.7e6, "c5b": 50.2e6}


def source_hash():
    """SHA-256 over the CUDA sources: ties an ncu capture (profiles/traffic.json) to the kernels it was taken from."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "vulkan-raytracing_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".h", ".hpp")):
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def roofline(st_count, trav_ms, frame_ms, hbm_gbs, hbm_src, fp32_tflops, npix, traffic_applies=True):
    """Algorithmic work of the traversal kernels (k_trace_shade + k_shadow) of one frame, DESIGN.md §5."""
    n_node, n_tri = st_count["nodes_visited"], st_count["tris_tested"]
    n_closest = st_count["rays_primary"] + st_count["rays_secondary"] - st_count.get("rays_top_level", 0)  # rays in the queue
    n_shadow = st_count["rays_shadow"]
    node_b, tri_b = st_count["node_bytes"], st_count["tri_bytes"]
    # ray state of the fused closest-hit + shading launch: 32 B read per traced ray, 32 B written per continuation ray, 48 B
    # per shadow ray appended; the shadow launch reads those 48 B and writes the 16 B radiance slot of an unoccluded one
    trav_bytes = (n_node * node_b + n_tri * tri_b + n_closest * 32 + st_count["rays_secondary"] * 32
                  + n_shadow * (48 + 48 + 16))
    children = {80: 8, 128: 4}.get(node_b, 2)      # boxes tested per node visit of the active layout (8-wide / 4-wide / binary)
    trav_flops = n_node * children * 24 + n_tri * 50
    achieved = trav_bytes / (trav_ms * 1e-3) / 1e9
    frame_bytes = (trav_bytes + st_count["n_hit"] * 84 + st_count["n_miss"] * 16 + n_closest * 32 + npix * 4)
    frame_flops = trav_flops + st_count["n_hit"] * 120
    t_roof = max(frame_bytes / (hbm_gbs * 1e9), frame_flops / (fp32_tflops * 1e12) if fp32_tflops else 0.0)
    # ncu figures of the traversal kernels (one --set full capture of this workload): only valid for the kernels they
    # were captured from, so the file carries the hash of the CUDA sources and everything from it is null otherwise
    path = os.path.join(ROOT, "profiles", "traffic.json")
    traffic = issue_frac = simd_lanes = None
    ncu_note = "no ncu capture applies to this workload"
    if traffic_applies and os.path.exists(path):
        try:
            t = json.load(open(path))
            if t.get("source_hash") == source_hash():
                traffic = t.get("k_trace_dram_bytes_per_frame")
                slots = trav_ms * 1e-3 * t["sm_count"] * 4 * t["sm_clock_hz"]   # issue slots of all SM sub-partitions in the traversal time
                issue_frac = round(t["warp_instructions_per_frame"] / slots, 4)
                simd_lanes = round(t["thread_instructions_per_frame"] / t["warp_instructions_per_frame"], 2)
                ncu_note = f"ncu capture {t.get('capture', '')} of these sources"
            else:
                ncu_note = "profiles/traffic.json was captured from other kernel sources (hash differs): traffic, issue_frac, simd_lanes withheld"
        except Exception as e:
            ncu_note = f"profiles/traffic.json unreadable: {e}"
    return {
        "bound": "hbm", "kernel": "k_trace_shade+k_shadow (all traversal launches of one frame)",
        "achieved": round(achieved, 1), "peak": hbm_gbs, "peak_source": hbm_src, "unit": "GB/s",
        "frac": round(achieved / hbm_gbs, 4), "traffic": traffic,
        "frac_algorithmic_cache_blind": round(achieved / hbm_gbs, 4),
        "issue_frac": issue_frac, "simd_lanes": simd_lanes, "ncu": ncu_note,
        "note": "achieved / frac follow SURVEY 8d: algorithmic node + triangle + ray-state bytes / traversal time with cache hits NOT "
                "discounted, so frac > 1 only says the working set is served from L1/L2 (`traffic` = what DRAM really moved, ncu). "
                "What binds the kernel: issue_frac = warp instructions / issue slots available in the traversal time, at simd_lanes "
                "active threads per warp instruction (DESIGN.md 5.3)",
        "algorithmic_bytes_per_frame": int(trav_bytes), "traversal_ms": round(trav_ms, 4),
        "nodes_visited": int(n_node), "tris_tested": int(n_tri), "node_bytes": node_b, "tri_bytes": tri_b, "boxes_per_node": children,
        "flop_term": {"flops_per_frame": int(trav_flops), "achieved_tflops": round(trav_flops / (trav_ms * 1e-3) / 1e12, 3),
                      "fp32_peak_tflops_measured": round(fp32_tflops, 2),
                      "frac": round(trav_flops / (trav_ms * 1e-3) / 1e12 / fp32_tflops, 4) if fp32_tflops else None},
        "frame_roofline_frac": round(t_roof / (frame_ms * 1e-3), 4),
    }


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = args.workload or ("c3" if args.gpus <= 1 else "c4")
    builder, W, H, depth, spp = WORKLOADS[wl]
    config = {"workload": WORKLOAD_TEXT[wl], "width": W, "height": H, "spp": spp, "max_bounce_count": depth,
              "skybox": f"procedural {args.sky}x{args.sky}x6 RGBA8" if wl != "c1" else "skybox_texture_sea fixture 256x256x6 RGBA8",
              "mesh": {"c2": "teapot", "c1": "cube_scene"}.get(wl, "armadillo_standin"),
              "l2": "ray-state buffers (GBs per frame) exceed L2; a 512 MiB buffer is also overwritten between steps"}

    if args.impl == "reference":
        if rank != 0:
            return 0
        graft.build() if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")) else None
        rt = graft.load_package()
        import scenes
        inputs = getattr(scenes, builder)(rt, sky_size=args.sky)
        params = rt.FrameParams.make(W, H, depth, spp)
        cam_ref = scenes.c5_camera(rt) if wl.startswith("c5") else rt.CameraVectors.default()
        v, ms, cores, sample, _, _ = cpu_arm(rt, inputs, cam_ref, params, args.steps, args.warmup, RAYS_PER_FRAME_HINT[wl])
        print(json.dumps({
            "impl": "reference", "metric": "Mrays/s", "value": round(v, 3), "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": round(v, 3), "unit": "Mrays/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": round(v, 3), "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "the reference has no CPU ray tracer and its Vulkan path is unavailable on this box; "
                    "this is the oracle port of its shaders on the host cores"}))
        return 0

    import torch
    import torch.distributed as dist
    rt = graft.load_package()
    import scenes
    if rt.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: the render path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    inputs = getattr(scenes, builder)(rt, sky_size=args.sky)
    scene = rt.Scene(inputs, device_id=local_rank, accel=args.accel | args.flags)
    build_ms = scene.stats()["build_ms"]
    scene.set_stream(torch.cuda.current_stream().cuda_stream)
    cam = scenes.c5_camera(rt) if wl.startswith("c5") else rt.CameraVectors.default()
    sharded = world > 1
    p_compact = rt.FrameParams.make(W, H, depth, spp, output_location=rt.LOCATION_DEVICE,
                                    shard_index=rank if sharded else 0, shard_count=world if sharded else 1)
    npix_local = int(rt.rt_lib().rt_frame_shard_pixels(C.byref(p_compact))) if sharded else W * H
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    launches = {"n": 0, "frames": 0}   # n: kernels around the frame (signal / wait / assemble); frames: frames rendered
    host = torch.zeros((H, W, 4), dtype=torch.uint8).pin_memory() if rank == 0 else None

    # ---- how the tiles reach rank 0 -----------------------------------------------------------------
    gather, peer = "none", None
    if sharded:
        gather = args.gather
        if gather == "peer":   # rank 0 owns the frame, everybody else maps it (CUDA IPC, NVLink peer stores)
            ok = torch.ones(1, device=dev)
            handle = torch.zeros(64, dtype=torch.uint8, device=dev)
            if rank == 0:
                try:
                    peer = rt.PeerFrame.create(local_rank, W * H * 4)
                    handle.copy_(torch.frombuffer(bytearray(peer.handle_bytes()), dtype=torch.uint8))
                except rt.RtError as e:
                    sys.stderr.write(f"peer frame: {e}\n")
                    ok.zero_()
            dist.broadcast(handle, src=0)
            if rank != 0:
                try:
                    peer = rt.PeerFrame.open(local_rank, handle.cpu().numpy().tobytes(), W * H * 4)
                except rt.RtError as e:
                    sys.stderr.write(f"rank {rank} peer frame: {e}\n")
                    ok.zero_()
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if float(ok.item()) == 0.0:
                if peer is not None:
                    peer.release()
                    peer = None
                gather = "nccl"
    if gather == "peer":
        p_dev = p_compact.copy()
        p_dev.shard_layout = rt.SHARD_IN_FRAME
        frame = torch.as_tensor(peer, device=dev).view(H, W, 4) if rank == 0 else None
        local = gathered = None
    else:
        p_dev = p_compact
        local = torch.zeros((npix_local, 4), dtype=torch.uint8, device=dev)
        gathered = torch.zeros((world, npix_local, 4), dtype=torch.uint8, device=dev) if (sharded and rank == 0) else None
        frame = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev) if (sharded and rank == 0) else None

    frame_no = {"n": 0}
    stream_handle = torch.cuda.current_stream().cuda_stream

    # ---- end-to-end path: one frame in shared pinned host memory, every rank stores its tiles into it --------------
    hostf, e2e_mode = None, args.e2e
    if e2e_mode == "host-frame":
        names = [f"/rt_b200_frame_{os.getpid()}"] if rank == 0 else [None]
        if world > 1:
            dist.broadcast_object_list(names, src=0)
        ok = torch.ones(1, device=dev)
        try:
            if rank == 0:
                hostf = rt.HostFrame.create(names[0], local_rank, W * H * 4)
        except rt.RtError as e:
            sys.stderr.write(f"host frame: {e}\n")
            ok.zero_()
        if world > 1:
            dist.barrier()          # the name exists before anybody opens it
        try:
            if rank != 0:
                hostf = rt.HostFrame.open(names[0], local_rank, W * H * 4)
        except rt.RtError as e:
            sys.stderr.write(f"rank {rank} host frame: {e}\n")
            ok.zero_()
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if float(ok.item()) == 0.0:
            if hostf is not None:
                hostf.release()
                hostf = None
            e2e_mode = "copy"
    p_host = p_compact.copy()
    p_host.shard_layout = rt.SHARD_IN_FRAME
    host_no = {"n": 0}

    def host_step():
        """one frame straight into the shared host frame: no device frame, no gather, no copy"""
        host_no["n"] += 1
        scene.render_into(cam, p_host, hostf.dev_ptr)    # returns when this rank's stores have completed
        launches["frames"] += 1
        hostf.signal(rank, host_no["n"])
        if rank == 0:
            hostf.wait(world, host_no["n"])

    def nccl_step(p, loc, gat, frm):
        scene.render_into(cam, p, loc.data_ptr())
        launches["frames"] += 1
        dist.gather(loc, [gat[i] for i in range(world)] if rank == 0 else None, dst=0)
        if rank == 0:
            scene.assemble_shards(p, gat.data_ptr(), frm.data_ptr())
            launches["n"] += 1

    def step(to_host):
        if to_host and e2e_mode == "host-frame":
            return host_step()
        if gather == "peer":
            frame_no["n"] += 1
            scene.render_into(cam, p_dev, peer.ptr)   # this rank's tiles go straight into rank 0's frame
            launches["frames"] += 1
            launches["n"] += 1
            peer.signal(rank, frame_no["n"], stream_handle)          # arrival flag in rank 0's memory, behind the stores
            if rank == 0:
                peer.wait(world, frame_no["n"], stream_handle)       # every rank's tiles are in place
                launches["n"] += 1
        elif gather == "nccl":
            nccl_step(p_dev, local, gathered, frame)
        else:
            scene.render_into(cam, p_dev, local.data_ptr())
            launches["frames"] += 1
        if to_host and rank == 0:
            src = frame if sharded else local.view(H, W, 4)
            host.copy_(src, non_blocking=True)
            torch.cuda.current_stream().synchronize()

    gather_verified = None
    if gather == "peer":   # once, outside the timed region: the collective path must give the same frame
        step(False)
        loc = torch.zeros((npix_local, 4), dtype=torch.uint8, device=dev)
        gat = torch.zeros((world, npix_local, 4), dtype=torch.uint8, device=dev) if rank == 0 else None
        frm = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev) if rank == 0 else None
        nccl_step(p_compact, loc, gat, frm)
        torch.cuda.synchronize()
        if rank == 0:
            gather_verified = bool(torch.equal(frm, frame))
        del loc, gat, frm
        torch.cuda.empty_cache()

    host_verified = None
    if e2e_mode == "host-frame":   # once, outside the timed region: the host frame must equal the device frame
        step(False)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        step(True)
        if rank == 0:
            dev_frame = (frame if sharded else local.view(H, W, 4)).cpu().numpy()
            host_verified = bool((hostf.array(H, W, 4) == dev_frame).all())

    def timed(to_host, steps):
        ms, rays = 0.0, 0
        for _ in range(steps):
            flush.zero_()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            step(to_host)
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
            st = scene.stats()
            r = torch.tensor([st["rays_primary"] + st["rays_secondary"] + st["rays_shadow"]], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dist.all_reduce(r, op=dist.ReduceOp.SUM)
            ms += float(t.item())
            rays += int(r.item())
        return ms, rays

    for _ in range(max(args.warmup, 3)):
        step(False)
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches["n"] = launches["frames"] = 0
    ms_dev, rays_dev = timed(False, args.steps)
    # rt_stats.kernel_launches of a frame (read outside the timed region: the library works the stats out on demand)
    n_launch = launches["n"] + launches["frames"] * scene.stats()["kernel_launches"]
    st_last = scene.stats()
    ms_e2e, rays_e2e = timed(True, args.steps)
    clocks = sampler.stop() if rank == 0 else None

    # counting frame (outside the timed region): exact node / triangle-test counts
    scene.set_counting(True)
    scene.render_into(cam, p_dev, peer.ptr if gather == "peer" else local.data_ptr())
    st_count = scene.stats()
    scene.set_counting(False)
    fp32 = rt.measure_fp32_tflops(local_rank) if rank == 0 else 0.0

    def solo_frames(p, ptr, n_warm, n_timed):
        """rank 0 alone, device output: mean frame ms (CUDA events) and rays per frame"""
        for _ in range(n_warm):
            scene.render_into(cam, p, ptr)
        ms = 0.0
        for _ in range(n_timed):
            flush.zero_()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            scene.render_into(cam, p, ptr)
            e1.record()
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
        st = scene.stats()
        return ms / n_timed, st["rays_primary"] + st["rays_secondary"] + st["rays_shadow"], st

    # (a) N > 1: the SAME workload on one GPU (rank 0 alone, the other ranks wait at the barrier), so that the line
    # carries its own strong-scaling denominator; N = 1 of the driver's scaling run is C3, a different frame
    solo = None
    if sharded:
        if rank == 0:
            full = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev)
            p_full = rt.FrameParams.make(W, H, depth, spp, output_location=rt.LOCATION_DEVICE)
            ms1, rays1, _ = solo_frames(p_full, full.data_ptr(), 1, 3)
            solo = {"ms_per_step": round(ms1, 4), "value": round(rays1 / (ms1 * 1e-3) / 1e6, 2), "unit": "Mrays/s", "frames": 3,
                    "equals_sharded_frame": bool(torch.equal(full, frame)) if frame is not None else None}
            del full
        dist.barrier()
    # (b) the rays the CPU arm renders (every k-th tile), on the GPU: same pixels, same rays
    same_rays = None
    if rank == 0 and not sharded:
        p_s, k_s = sample_params(rt, p_compact, st_last["rays_primary"] + st_last["rays_secondary"] + st_last["rays_shadow"])
        p_s.output_location = rt.LOCATION_DEVICE
        buf = torch.zeros((int(rt.rt_lib().rt_frame_shard_pixels(C.byref(p_s))), 4), dtype=torch.uint8, device=dev)
        ms_s, rays_s, _ = solo_frames(p_s, buf.data_ptr(), 2, 5)
        same_rays = {"tiles": f"every {k_s}th 64x64 tile", "rays": int(rays_s), "gpu_ms": round(ms_s, 4),
                     "gpu_value": round(rays_s / (ms_s * 1e-3) / 1e6, 2), "unit": "Mrays/s", "_image": buf.cpu().numpy()}
        del buf
    # (c) the drop-in call of INTEGRATION.md: rt_render_frame(RT_LOCATION_HOST) into a plain malloc'd buffer, as is
    # (device frame + pageable copy) and after rt_host_buffer_pin (resolve kernel stores straight into it)
    plain = None
    if rank == 0 and not sharded:
        p_h = rt.FrameParams.make(W, H, depth, spp)   # RT_LOCATION_HOST, full frame
        hostbuf = np.empty((H, W, 4), np.uint8)
        res = {}
        for mode in ("pageable_copy", "pinned_by_rt_host_buffer_pin"):
            if mode != "pageable_copy":
                scene.pin_host_buffer(hostbuf)
            for _ in range(2):
                scene.render_into(cam, p_h, hostbuf.ctypes.data)
            t_ms = 0.0
            for _ in range(5):
                flush.zero_()
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                scene.render_into(cam, p_h, hostbuf.ctypes.data)   # synchronous: the pixels are in hostbuf on return
                t_ms += 1e3 * (time.perf_counter() - t0)
            st_h = scene.stats()
            rays_h = st_h["rays_primary"] + st_h["rays_secondary"] + st_h["rays_shadow"]
            res[mode] = {"ms_per_step": round(t_ms / 5, 4), "value": round(rays_h / (t_ms / 5 * 1e-3) / 1e6, 2)}
            res[mode]["equals_device_frame"] = bool((hostbuf == local.view(H, W, 4).cpu().numpy()).all())
        scene.unpin_host_buffer()
        plain = {"unit": "Mrays/s", "timer": "host wall clock around the synchronous call", **res}
    if rank != 0:
        if hostf is not None:
            torch.cuda.synchronize()
            hostf.release()
        if peer is not None:
            torch.cuda.synchronize()
            peer.release()          # peers unmap before the owner frees
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    hbm, hbm_src = load_peaks()
    value = rays_dev / (ms_dev * 1e-3) / 1e6
    e2e_v = rays_e2e / (ms_e2e * 1e-3) / 1e6
    out = {
        "metric": "Mrays/s", "value": round(value, 2), "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(ms_dev / args.steps, 4), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
        "frame_ms": round(ms_dev / args.steps, 4), "rays_per_frame": rays_dev // args.steps, "build_ms": round(build_ms, 3),
        "e2e": {"value": round(e2e_v, 2), "unit": "Mrays/s", "ms_per_step": round(ms_e2e / args.steps, 4),
                "h2d_bytes_per_step": C.sizeof(rt.CameraVectors) + C.sizeof(rt.FrameParams), "d2h_bytes_per_step": W * H * 4},
        "gpu_launches": int(n_launch), "clocks": clocks,
        "e2e_path": {"host-frame": "k_resolve of every rank stores its tiles into one frame in shared pinned host memory "
                                   "(rt_host_frame_*, 128-byte runs over PCIe) + one CPU arrival flag per rank",
                     "copy": "frame left on rank 0's GPU, then one device->pinned-host copy"}[e2e_mode],
        "stage_ms": {"trace": round(st_last["trace_ms"], 4), "shadow": round(st_last["shadow_ms"], 4),
                     "shade_raygen_resolve": round(st_last["shade_ms"], 4), "frame_in_library": round(st_last["frame_ms"], 4)},
        "bvh": {"nodes": st_last["bvh_nodes"], "bytes": st_last["bvh_bytes"], "node_bytes": st_last["node_bytes"],
                "layout": {64: "binary, 2 FP32 boxes per 64 B node", 80: "8-wide, quantised, 80 B node",
                           128: "4-wide, 4 FP32 boxes per 128 B node"}.get(st_last["node_bytes"], "none"),
                "depth": st_last["max_tree_depth"]},
    }
    if sharded:
        out["gather"] = {
            "peer": "k_resolve of every rank stores its tiles into rank 0's frame (CUDA IPC mapping, NVLink peer stores), "
                    "then one arrival flag per rank that rank 0 waits for; no collective, no de-interleave pass",
            "nccl": "torch.distributed.gather (NCCL) of the compact tile buffers + rt_assemble_shards"}[gather]
        out["gather_mode"] = gather
        if gather_verified is not None:
            out["gather_equals_nccl_path"] = gather_verified
    if host_verified is not None:
        out["host_frame_equals_device_frame"] = host_verified
    out["roofline"] = roofline(st_count, st_last["trace_ms"] + st_last["shadow_ms"], st_last["frame_ms"], hbm, hbm_src, fp32,
                               npix_local, traffic_applies=(wl == "c3" and world == 1 and args.accel in (0, 1)))
    out["rays"] = {"primary": st_last["rays_primary"], "of_which_answered_by_the_top_level_in_raygen": st_last["rays_top_level"],
                   "secondary": st_last["rays_secondary"], "shadow": st_last["rays_shadow"],
                   "note": "value counts every traceRayEXT-equivalent, as the oracle does; primary rays that reach no instance box are "
                           "answered in k_raygen (sky lookup) and never traversed"}
    trav_rays = out["rays_per_frame"] - st_last["rays_top_level"]
    out["value_traversed_rays_only"] = round(trav_rays / (ms_dev / args.steps * 1e-3) / 1e6, 2)
    if solo is not None:
        out["single_gpu_same_workload"] = solo
        out["strong_efficiency_same_workload"] = round(value / (world * solo["value"]), 4)
        out["data_plane"] = ("tiles: peer stores into rank 0's frame + arrival flags (no collective)" if gather == "peer" else
                             "tiles: NCCL gather + de-interleave kernel") + "; NCCL otherwise carries only the set-up broadcasts and the bench barriers"
    if plain is not None:
        out["e2e_plain_host"] = plain
    if world == 1 and not args.no_cpu_baseline:
        p_cpu = rt.FrameParams.make(W, H, depth, spp)
        v, ms, cores, sample, rays_c, img_c = cpu_arm(rt, inputs, cam, p_cpu, 1, 0, out["rays_per_frame"], want_image=True)
        out["cpu_baseline"] = {"value": round(v, 3), "unit": "Mrays/s", "cores": cores, "kind": "port", "sample": sample,
                               "ms_per_sample": round(ms, 1)}
        if same_rays is not None:
            same_rays["cpu_value"] = round(v, 3)
            same_rays["gpu_over_cpu"] = round(same_rays["gpu_value"] / v, 1) if v else None
            same_rays["rays_equal"] = bool(rays_c == same_rays["rays"])
            same_rays["images_equal"] = bool((img_c == same_rays["_image"]).all())
    if same_rays is not None:
        same_rays.pop("_image", None)
        out["same_rays"] = same_rays
    print(json.dumps(out))
    if hostf is not None:
        hostf.release()             # the creator also removes the name
    if world > 1:
        dist.barrier()              # the peers have unmapped the frame
        if peer is not None:
            del frame
            peer.release()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
