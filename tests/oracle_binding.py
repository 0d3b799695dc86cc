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
