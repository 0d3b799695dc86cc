import sys, os, ctypes as C
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "tests"))
import torch
import __graft_entry__ as graft
rt = graft.load_package()
import scenes
inputs = scenes.c3_scene(rt, sky_size=2048)
scene = rt.Scene(inputs, device_id=0)
scene.set_stream(torch.cuda.current_stream().cuda_stream)
cam = rt.CameraVectors.default()
W, H, idx, cnt = (int(a) for a in sys.argv[1:5])
p = rt.FrameParams.make(W, H, 8, 4, output_location=rt.LOCATION_DEVICE, shard_index=idx, shard_count=cnt)
npix = int(rt.rt_lib().rt_frame_shard_pixels(C.byref(p))) if cnt > 1 else W * H
out = torch.zeros((npix, 4), dtype=torch.uint8, device="cuda")
for _ in range(4):
    sys.stderr.write("---- frame\n"); sys.stderr.flush()
    scene.render_into(cam, p, out.data_ptr())
    torch.cuda.synchronize()
st = scene.stats()
sys.stderr.write(f"frame {st['frame_ms']:.3f} trace {st['trace_ms']:.3f} shadow {st['shadow_ms']:.3f} shade {st['shade_ms']:.3f}\n")
