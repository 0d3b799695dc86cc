"""ctypes binding of tests/emu/libemu.so: the product's RT_HD device functions (LBVH build,
compressed-wide traversal, shading) compiled for the CPU and run serially.  Tests only."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(os.path.join(ROOT, "tests", "emu", "libemu.so"))
        L.emu_scene_create.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.emu_scene_destroy.argtypes = [C.c_void_p]
        L.emu_scene_set_materials.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.emu_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_void_p]
        L.emu_trace.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_int, C.c_void_p]
        L.emu_bvh_nodes.restype = C.c_uint64
        L.emu_bvh_nodes.argtypes = [C.c_void_p, C.c_uint32]
        L.emu_wide_nodes.restype = C.c_void_p
        L.emu_wide_nodes.argtypes = [C.c_void_p, C.c_uint32]
        L.emu_bvh2_nodes.restype = C.c_void_p
        L.emu_bvh2_nodes.argtypes = [C.c_void_p, C.c_uint32]
        L.emu_leaf_tris.restype = C.c_void_p
        L.emu_leaf_tris.argtypes = [C.c_void_p, C.c_uint32]
        L.emu_treelet.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float),
                                  C.c_void_p, C.c_void_p]
        L.emu_slot_to_pixel.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        _lib = L
    return _lib


class EmuScene:
    def __init__(self, inputs, accel=0):
        self.inputs = inputs
        self._h = C.c_void_p()
        d = inputs.desc(flags=accel)
        assert lib().emu_scene_create(C.byref(d), C.byref(self._h)) == 0

    def __del__(self):
        if self._h and self._h.value:
            lib().emu_scene_destroy(self._h)
            self._h = C.c_void_p()

    def set_materials(self, materials):
        from vulkan_raytracing_b200 import material_array
        if materials is None:
            assert lib().emu_scene_set_materials(self._h, None, 0) == 0
        else:
            arr = material_array(materials)
            assert lib().emu_scene_set_materials(self._h, C.cast(arr, C.c_void_p), len(materials)) == 0

    def render(self, camera, params, crop=None):
        from vulkan_raytracing_b200 import HIT_DTYPE
        x0, y0, x1, y1 = crop if crop else (0, 0, params.width, params.height)
        w, h, spp = x1 - x0, y1 - y0, params.samples_per_pixel
        f32 = np.zeros((h, w, 4), np.float32)
        u8 = np.zeros((h, w, 4), np.uint8)
        prim = np.zeros(h * w * spp, HIT_DTYPE)
        cnt = np.zeros(8, np.uint64)
        lib().emu_render(self._h, C.byref(camera), C.byref(params), x0, y0, x1, y1, f32.ctypes.data, u8.ctypes.data,
                         prim.ctypes.data, cnt.ctypes.data)
        names = ["primary", "secondary", "shadow", "miss", "hit", "_", "nodes", "tris"]
        return {"rgba32f": f32, "rgba8": u8, "primary": prim, "counts": {n: int(c) for n, c in zip(names, cnt) if n != "_"}}

    def trace_rays(self, origins, dirs, tmin=0.001, tmax=10000.0, any_hit=False):
        from vulkan_raytracing_b200 import HIT_DTYPE
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        out = np.zeros(o.shape[0], HIT_DTYPE)
        lib().emu_trace(self._h, o.shape[0], o.ctypes.data, d.ctypes.data, tmin, tmax, int(any_hit), out.ctypes.data)
        return out

    def wide_nodes(self, mesh=0):
        n = int(lib().emu_bvh_nodes(self._h, mesh))
        p = lib().emu_wide_nodes(self._h, mesh)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), (n, 20)).copy()

    def bvh2_nodes(self, mesh=0):
        """(n, 16) float32 view of the binary nodes: child boxes + child ids (as int bits)."""
        n = int(lib().emu_bvh_nodes(self._h, mesh))
        p = lib().emu_bvh2_nodes(self._h, mesh)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), (n, 16)).copy()

    def leaf_tris(self, mesh, ntri):
        p = lib().emu_leaf_tris(self._h, mesh)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), (ntri, 12)).copy()


def slot_to_pixel(params, pslot):
    x, y = C.c_uint32(), C.c_uint32()
    ok = lib().emu_slot_to_pixel(C.byref(params), pslot, C.byref(x), C.byref(y))
    return bool(ok), x.value, y.value


def slab_tests(o, d, box, tmin, tmax):
    """The node tests of rt_trav.cuh on explicit (ray, box) pairs: uint8 per pair, bit 0 = round-1 test, bit 1 = the
    default one-fma test, bit 2 = the sign-select variant."""
    o = np.ascontiguousarray(o, np.float32); d = np.ascontiguousarray(d, np.float32)
    box = np.ascontiguousarray(box, np.float32); tmax = np.ascontiguousarray(tmax, np.float32)
    n = o.shape[0]
    out = np.zeros(n, np.uint8)
    f = lib().emu_slab_tests
    f.argtypes = [C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p]
    assert f(n, o.ctypes.data, d.ctypes.data, box.ctypes.data, float(tmin), tmax.ctypes.data, out.ctypes.data) == 0
    return out


def treelet(boxes, leaf_cost):
    """Device treelet optimisation of one root over n <= 7 leaves (emu_treelet): returns
    (cost of the initial chain, cost after, left[n-1], right[n-1])."""
    b = np.ascontiguousarray(boxes, np.float32).reshape(-1, 6)
    c = np.ascontiguousarray(leaf_cost, np.float32)
    n = b.shape[0]
    before, after = C.c_float(), C.c_float()
    left, right = np.zeros(n - 1, np.int32), np.zeros(n - 1, np.int32)
    assert lib().emu_treelet(n, b.ctypes.data, c.ctypes.data, C.byref(before), C.byref(after), left.ctypes.data,
                             right.ctypes.data) == 0
    return before.value, after.value, left, right
