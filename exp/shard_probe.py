# one GPU: time a C4 shard (k of n) against the full C3 frame, same ray count at n = 4
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import __graft_entry__ as graft
rt = graft.load_package()
import scenes
inputs = scenes.c3_scene(rt, sky_size=2048)
scene = rt.Scene(inputs, device_id=0)
scene.set_stream(torch.cuda.current_stream().cuda_stream)
cam = rt.CameraVectors.default()
def run(W, H, idx, cnt, label):
    p = rt.FrameParams.make(W, H, 8, 4, output_location=rt.LOCATION_DEVICE, shard_index=idx, shard_count=cnt)
    npix = int(rt.rt_lib().rt_frame_shard_pixels(C.byref(p))) if cnt > 1 else W * H
    out = torch.zeros((npix, 4), dtype=torch.uint8, device="cuda")
    for _ in range(3):
        scene.render_into(cam, p, out.data_ptr())
    torch.cuda.synchronize()
    acc = {}
    for _ in range(5):
        scene.render_into(cam, p, out.data_ptr())
        torch.cuda.synchronize()
        st = scene.stats()
        for k in ("frame_ms", "trace_ms", "shadow_ms", "shade_ms"):
            acc[k] = acc.get(k, 0.0) + st[k] / 5
    rays = st["rays_primary"] + st["rays_secondary"] + st["rays_shadow"]
    print(f"{label:24s} rays {rays/1e6:7.2f}M frame {acc['frame_ms']:7.3f} trace {acc['trace_ms']:7.3f} shadow {acc['shadow_ms']:6.3f} shade {acc['shade_ms']:6.3f}  Mrays/s {rays/acc['frame_ms']/1e3:8.1f}", flush=True)
run(3840, 2160, 0, 1, "C3 full")
for n in ((8,) if os.environ.get("PROBE_SHORT") else (2, 4, 8)):
    for k in sorted({0} if os.environ.get("PROBE_SHORT") else {0, n - 1}):
        run(7680, 4320, k, n, f"C4 shard {k}/{n}")
run(7680, 4320, 0, 1, "C4 full")
