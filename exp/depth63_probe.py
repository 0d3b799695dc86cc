# C3's scene at the reference's default MAX_BOUNCE_COUNT 63 (config.h:26), 3840x2160: frame time with the path-kernel tail (RT_TAIL_FROM, read
# once per process) - usage: RT_TAIL_FROM=9|0 python exp/depth63_probe.py
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import __graft_entry__ as graft
rt = graft.load_package()
import scenes
scene = rt.Scene(scenes.c3_scene(rt, sky_size=2048), device_id=0)
cam = rt.CameraVectors.default()
p = rt.FrameParams.make(3840, 2160, 63, 4, output_location=rt.LOCATION_DEVICE)
out = torch.zeros((3840 * 2160, 4), dtype=torch.uint8, device="cuda")
for _ in range(3):
    scene.render_into(cam, p, out.data_ptr())
torch.cuda.synchronize()
acc = 0.0
for _ in range(5):
    scene.render_into(cam, p, out.data_ptr())
    torch.cuda.synchronize()
    st = scene.stats()
    acc += st["frame_ms"] / 5
rays = st["rays_primary"] + st["rays_secondary"] + st["rays_shadow"]
print(f"C3 scene, depth 63, RT_TAIL_FROM={os.environ.get('RT_TAIL_FROM', '9 (default)')}: frame {acc:.3f} ms, {rays/1e6:.2f} M rays, {st['kernel_launches']} launches, "
      f"sha {__import__('hashlib').sha256(out.cpu().numpy().tobytes()).hexdigest()[:16]}", flush=True)
