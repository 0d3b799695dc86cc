"""ctypes binding of oracle/liboracle.so — the CPU checker. Test infrastructure only."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_LIB = os.path.join(ROOT, "oracle", "liboracle.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(ORACLE_LIB)
        L.orc_last_error.restype = C.c_char_p
        L.orc_scene_create.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.orc_scene_destroy.argtypes = [C.c_void_p]
        L.orc_scene_set_materials.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.orc_scene_set_transforms.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.orc_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                 C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_render_shard.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_trace.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_int,
                                C.c_int, C.c_void_p]
        L.orc_trace_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double,
                                    C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_double)]
        L.orc_det_sin.restype = C.c_float
        L.orc_det_powf.restype = C.c_float
        L.orc_det_powf.argtypes = [C.c_float, C.c_float]
        L.orc_det_sin.argtypes = [C.c_float]
        L.orc_random.restype = C.c_float
        L.orc_random.argtypes = [C.c_float, C.c_float, C.c_float]
        L.orc_sky_sample.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_invert34.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_bvh_nodes.restype = C.c_uint64
        L.orc_bvh_nodes.argtypes = [C.c_void_p, C.c_uint32]
        _lib = L
    return _lib


COUNT_NAMES = ["primary", "secondary", "shadow", "miss", "hit", "tri_tests"]


class OracleScene:
    """Same inputs as the product's Scene (a SceneInputs), rendered by the CPU oracle."""

    def __init__(self, inputs):
        self.inputs = inputs
        self._h = C.c_void_p()
        d = inputs.desc()
        rc = lib().orc_scene_create(C.byref(d), C.byref(self._h))
        if rc != 0:
            raise RuntimeError(f"oracle error {rc}: {lib().orc_last_error().decode()}")

    def close(self):
        if self._h and self._h.value:
            lib().orc_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_transforms(self, xforms):
        a = np.ascontiguousarray(xforms, np.float32).reshape(-1, 12)
        assert lib().orc_scene_set_transforms(self._h, a.ctypes.data, a.shape[0]) == 0

    def set_materials(self, materials):
        from vulkan_raytracing_b200 import material_array
        if materials is None:
            assert lib().orc_scene_set_materials(self._h, None, 0) == 0
        else:
            arr = material_array(materials)
            assert lib().orc_scene_set_materials(self._h, C.cast(arr, C.c_void_p), len(materials)) == 0

    def render(self, camera, params, crop=None, brute=False, threads=0, want=("rgba8",)):
        """Returns a dict with any of rgba32f, rgba8, primary, counts over the crop (x0,y0,x1,y1)."""
        from vulkan_raytracing_b200 import HIT_DTYPE
        x0, y0, x1, y1 = crop if crop else (0, 0, params.width, params.height)
        w, h, spp = x1 - x0, y1 - y0, params.samples_per_pixel
        f32 = np.zeros((h, w, 4), np.float32) if "rgba32f" in want else None
        u8 = np.zeros((h, w, 4), np.uint8) if "rgba8" in want else None
        prim = np.zeros(h * w * spp, HIT_DTYPE) if "primary" in want else None
        cnt = np.zeros(8, np.uint64) if "counts" in want else None
        ptr = lambda a: a.ctypes.data if a is not None else None
        rc = lib().orc_render(self._h, C.byref(camera), C.byref(params), x0, y0, x1, y1, int(brute), threads,
                              ptr(f32), ptr(u8), ptr(prim), ptr(cnt))
        if rc != 0:
            raise RuntimeError(f"oracle render error {rc}")
        out = {"rgba32f": f32, "rgba8": u8, "primary": prim}
        if cnt is not None:
            out["counts"] = {n: int(cnt[i]) for i, n in enumerate(COUNT_NAMES)}
        return {k: v for k, v in out.items() if v is not None}

    def render_shard(self, camera, params, threads=0, want_image=True):
        """The tiles of shard params.shard_index of params.shard_count in the C ABI's compact shard layout
        (uint8[n_pixels, 4]) plus the ray counts."""
        tw, th = params.tile_w or 64, params.tile_h or 64
        tiles = ((params.width + tw - 1) // tw) * ((params.height + th - 1) // th)
        n = max(1, params.shard_count)
        per = (tiles + n - 1) // n * tw * th
        img = np.zeros((per, 4), np.uint8) if want_image else None
        cnt = np.zeros(8, np.uint64)
        rc = lib().orc_render_shard(self._h, C.byref(camera), C.byref(params), 0, threads,
                                    img.ctypes.data if img is not None else None, cnt.ctypes.data)
        if rc != 0:
            raise RuntimeError(f"oracle render_shard error {rc}")
        return img, {k: int(cnt[i]) for i, k in enumerate(COUNT_NAMES)}

    def trace_rays(self, origins, dirs, tmin=0.001, tmax=10000.0, any_hit=False, brute=False):
        from vulkan_raytracing_b200 import HIT_DTYPE
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        out = np.zeros(o.shape[0], HIT_DTYPE)
        assert lib().orc_trace(self._h, o.shape[0], o.ctypes.data, d.ctypes.data, tmin, tmax, int(any_hit),
                               int(brute), out.ctypes.data) == 0
        return out

    def trace_f64(self, origin, direction, tmin=0.001, tmax=10000.0):
        o = np.asarray(origin, np.float32)
        d = np.asarray(direction, np.float32)
        i, p, t = C.c_uint32(), C.c_uint32(), C.c_double()
        lib().orc_trace_f64(self._h, o.ctypes.data, d.ctypes.data, tmin, tmax, C.byref(i), C.byref(p), C.byref(t))
        return i.value, p.value, t.value

    def sky_sample(self, direction):
        d = np.asarray(direction, np.float32)
        out = np.zeros(3, np.float32)
        lib().orc_sky_sample(self._h, d.ctypes.data, out.ctypes.data)
        return out


def det_powf(x, y):
    return lib().orc_det_powf(float(x), float(y))


def det_sin(x):
    return lib().orc_det_sin(float(x))


def random_hash(ux, uy, seed):
    return lib().orc_random(float(ux), float(uy), float(seed))


def invert34(m):
    a = np.ascontiguousarray(m, np.float32).reshape(12)
    out = np.zeros(12, np.float32)
    lib().orc_invert34(a.ctypes.data, out.ctypes.data)
    return out


# ----------------------------------------------------------------------------------------------
# oracle/liboracle64.so: the independent float64 renderer (oracle64.cpp) and the helpers that
# compare a frame with it at BASELINE.json's thresholds
ORACLE64_LIB = os.path.join(ROOT, "oracle", "liboracle64.so")
_lib64 = None


def lib64():
    global _lib64
    if _lib64 is None:
        L = C.CDLL(ORACLE64_LIB)
        L.o64_scene_create.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.o64_scene_destroy.argtypes = [C.c_void_p]
        L.o64_scene_set_transforms.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.o64_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.o64_trace.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_void_p,
                                C.c_void_p, C.c_void_p]
        lib().orc_jitter.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        lib().orc_render_sigs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.c_void_p]
        _lib64 = L
    return _lib64


def jitter(params, crop=None):
    """Pixel jitter of shader.rgen:72-73 under the FP32 contract: (h, w, spp, 2) float32 over the crop."""
    lib64()
    x0, y0, x1, y1 = crop if crop else (0, 0, params.width, params.height)
    out = np.zeros((y1 - y0, x1 - x0, params.samples_per_pixel, 2), np.float32)
    lib().orc_jitter(C.byref(params), x0, y0, x1, y1, out.ctypes.data)
    return out


def oracle32_signatures(scene32, camera, params, crop=None, brute=False):
    lib64()
    x0, y0, x1, y1 = crop if crop else (0, 0, params.width, params.height)
    sig = np.zeros((y1 - y0, x1 - x0, params.samples_per_pixel), np.uint64)
    assert lib().orc_render_sigs(scene32._h, C.byref(camera), C.byref(params), x0, y0, x1, y1, int(brute), sig.ctypes.data) == 0
    return sig


class Oracle64Scene:
    """Same inputs as the product's Scene, rendered by the independent float64 renderer."""

    def __init__(self, inputs):
        self.inputs = inputs
        self._h = C.c_void_p()
        d = inputs.desc()
        rc = lib64().o64_scene_create(C.byref(d), C.byref(self._h))
        if rc != 0:
            raise RuntimeError(f"oracle64 error {rc}")

    def close(self):
        if self._h and self._h.value:
            lib64().o64_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_transforms(self, xforms):
        a = np.ascontiguousarray(xforms, np.float32).reshape(-1, 12)
        assert lib64().o64_scene_set_transforms(self._h, a.ctypes.data, a.shape[0]) == 0

    def render(self, camera, params, crop=None, jitter_shift=0.0):
        from vulkan_raytracing_b200 import HIT_DTYPE
        x0, y0, x1, y1 = crop if crop else (0, 0, params.width, params.height)
        w, h, spp = x1 - x0, y1 - y0, params.samples_per_pixel
        jit = jitter(params, (x0, y0, x1, y1))
        if jitter_shift:
            jit = np.clip(jit + np.float32(jitter_shift), 0.0, 0.999999).astype(np.float32)
        f64 = np.zeros((h, w, 4), np.float64)
        u8 = np.zeros((h, w, 4), np.uint8)
        prim = np.zeros(h * w * spp, HIT_DTYPE)
        sig = np.zeros((h, w, spp), np.uint64)
        cnt = np.zeros(8, np.uint64)
        rc = lib64().o64_render(self._h, C.byref(camera), C.byref(params), x0, y0, x1, y1, jit.ctypes.data, f64.ctypes.data,
                                u8.ctypes.data, prim.ctypes.data, sig.ctypes.data, cnt.ctypes.data)
        if rc != 0:
            raise RuntimeError(f"oracle64 render error {rc}")
        return {"rgba64": f64, "rgba8": u8, "primary": prim, "sig": sig,
                "counts": {n: int(cnt[i]) for i, n in enumerate(COUNT_NAMES[:5])}}

    def sensitivity(self, camera, params, crop=None, base=None, shift=1e-3):
        """Per pixel: the largest 8-bit change of oracle64's own image when every sample's jitter moves by
        +-shift pixel (about the size of FP32 rounding of a hit position seen from the camera).  A pixel that
        moves by more than 2/255 under that is ill-conditioned: no two arithmetics can agree on it."""
        base = base if base is not None else self.render(camera, params, crop)
        sens = np.zeros(base["rgba8"].shape[:2], np.int32)
        for d in (shift, -shift):
            moved = self.render(camera, params, crop, jitter_shift=d)
            sens = np.maximum(sens, np.abs(moved["rgba8"].astype(np.int32) - base["rgba8"].astype(np.int32)).max(axis=2))
        return sens

    def trace_rays(self, origins, dirs, tmin=0.001, tmax=10000.0):
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        inst = np.zeros(o.shape[0], np.uint32)
        prim = np.zeros(o.shape[0], np.uint32)
        t = np.zeros(o.shape[0], np.float64)
        assert lib64().o64_trace(self._h, o.shape[0], o.ctypes.data, d.ctypes.data, float(np.float32(tmin)),
                                 float(np.float32(tmax)), inst.ctypes.data, prim.ctypes.data, t.ctypes.data) == 0
        return inst, prim, t


def compare_with_oracle64(img8, primary, ref64, sig32=None):
    """Frame (RGBA8 + primary hit records) against the float64 renderer at BASELINE.json's thresholds.
    Returns a report dict: primary-id match rate, max per-channel error, PSNR, and the residual pixels
    (error > 2/255) classified by cause:
      primary    a primary ray of the pixel hit another primitive (an edge / silhouette sample),
      path       same primary hits, but a later bounce of some sample took another path (sig32 given:
                 the two renderers' path signatures differ - chaotic amplification through the
                 refractive mesh), unknown when no signatures are available,
      arithmetic same paths, different value: that would be a formula disagreement."""
    spp = ref64["sig"].shape[2]
    h, w = ref64["rgba8"].shape[:2]
    rp = ref64["primary"]
    same = (primary["instance"] == rp["instance"]) & (primary["primitive"] == rp["primitive"])
    diff = np.abs(img8.astype(np.int32) - ref64["rgba8"].astype(np.int32)).max(axis=2)
    mse = np.mean((img8[..., :3].astype(np.float64) - ref64["rgba8"][..., :3].astype(np.float64)) ** 2)
    psnr = 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)
    resid = diff > 2
    prim_pix = ~same.reshape(h, w, spp).all(axis=2)
    rep = {"primary_match": float(same.mean()), "max_err": int(diff.max()), "psnr": float(psnr),
           "residual_pixels": int(resid.sum()), "pixels": int(h * w),
           "residual_primary": int((resid & prim_pix).sum())}
    if sig32 is not None:
        path_pix = (sig32 != ref64["sig"]).any(axis=2)
        rep["samples_path_diverged"] = int((sig32 != ref64["sig"]).sum())
        rep["residual_path"] = int((resid & ~prim_pix & path_pix).sum())
        rep["residual_arithmetic"] = int((resid & ~prim_pix & ~path_pix).sum())
        # error of the pixels whose every sample followed the same path in both renderers
        clean = ~path_pix & ~prim_pix
        rep["max_err_same_path"] = int(diff[clean].max()) if clean.any() else 0
        rep["pixels_same_path"] = int(clean.sum())
        rep["_arith_mask"] = resid & clean
    return rep


def explain_arithmetic_residuals(rep, scene64, camera, params, crop, ref64):
    """Splits rep['residual_arithmetic'] (same path in both renderers, error > 2/255) into ill-conditioned pixels
    (oracle64 itself moves by more than 2/255 under a 1e-3 pixel jitter shift: a refraction near the critical angle,
    a reflection chain grazing an edge, a sharp skybox texel seen through several bounces) and unexplained ones -
    the latter would be a disagreement between the two readings of the shaders.  Returns the unexplained count."""
    mask = rep.pop("_arith_mask", None)
    if mask is None or not mask.any():
        rep["residual_ill_conditioned"], rep["residual_unexplained"] = 0, 0
        return 0
    sens = scene64.sensitivity(camera, params, crop, base=ref64)
    rep["residual_ill_conditioned"] = int((mask & (sens > 2)).sum())
    rep["residual_unexplained"] = int((mask & (sens <= 2)).sum())
    return rep["residual_unexplained"]
