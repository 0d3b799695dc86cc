"""vulkan-raytracing_b200 — Python host mirror over the C ABI of the B200 ray-tracing core.

The product is ``librt_b200.so`` (hand-written CUDA for sm_100a behind ``include/rt_b200.h``)
plus ``librt_host.so`` (asset loading / camera, ``include/rt_host.h``).  This module only
binds them with ctypes for tests and ``bench.py``; it contains no rendering code and has
no CPU fallback: if the CUDA library is missing or no device is visible, every compute
call raises.

Because the directory name carries a hyphen, import it through
``__graft_entry__.load_package()`` (registers it as ``vulkan_raytracing_b200``).

Reference (``/root/reference``) anchors: ``Config`` = include/config.h:4-27;
``Camera`` = include/camera.h:16-35; ``Scene`` = src/main.cpp:1603-1835 + 2061-2412;
``Scene.render`` = vkCmdTraceRaysKHR, src/main.cpp:2620-2624.
"""
from __future__ import annotations

import ctypes as C
import dataclasses
import os
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_HERE)
RT_LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "librt_b200.so")
HOST_LIB_PATH = os.path.join(_HERE, "librt_host.so")

TYPE_DIFFUSE, TYPE_MIRROR, TYPE_REFRACTIVE = 0, 1, 2
FORMAT_RGBA8, FORMAT_RGBA32F = 0, 1
LOCATION_HOST, LOCATION_DEVICE = 0, 1
SHARD_COMPACT, SHARD_IN_FRAME = 0, 1
ACCEL_DEFAULT, ACCEL_BVH2, ACCEL_BRUTE, ACCEL_WIDE8, ACCEL_WIDE4 = 0, 1, 2, 3, 4
FLAG_NO_REFINE = 0x10
FLAG_GENERATE_NORMALS = 0x20
FLAG_SKY_TEXTURE = 0x40
SORT_NOALLOC_MAX_PAIRS = 1 << 20
MISS = 0xFFFFFFFF


ERR_INVALID_ARGUMENT, ERR_NO_DEVICE, ERR_CUDA, ERR_OUT_OF_MEMORY, ERR_UNSUPPORTED = -1, -2, -3, -4, -5


class RtError(RuntimeError):
    """Failure of a C-ABI call; `code` is the RT_ERR_* value (None for errors raised on the Python side)."""
    code = None


# ----------------------------------------------------------------------------- structs
class MeshDesc(C.Structure):
    _fields_ = [("vertices", C.POINTER(C.c_float)), ("n_vertices", C.c_uint32),
                ("indices", C.POINTER(C.c_uint32)), ("n_triangles", C.c_uint32)]


class InstanceDesc(C.Structure):
    _fields_ = [("mesh", C.c_uint32), ("type", C.c_uint32), ("transform", C.c_float * 12)]


class SceneDesc(C.Structure):
    _fields_ = [("n_meshes", C.c_uint32), ("meshes", C.POINTER(MeshDesc)),
                ("n_instances", C.c_uint32), ("instances", C.POINTER(InstanceDesc)),
                ("skybox_rgba8", C.POINTER(C.c_uint8)), ("sky_w", C.c_uint32), ("sky_h", C.c_uint32),
                ("device_id", C.c_int32), ("flags", C.c_uint32)]


class CameraVectors(C.Structure):
    """rt_camera: the four vectors of the uniform block (src/main.cpp:1848-1851)."""
    _fields_ = [("position", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3),
                ("forward", C.c_float * 3)]

    @staticmethod
    def default() -> "CameraVectors":
        return CameraVectors.make((0, 0, 20), (1, 0, 0), (0, 1, 0), (0, 0, -1))

    @staticmethod
    def make(position, right, up, forward) -> "CameraVectors":
        c = CameraVectors()
        c.position[:] = [float(x) for x in position]
        c.right[:] = [float(x) for x in right]
        c.up[:] = [float(x) for x in up]
        c.forward[:] = [float(x) for x in forward]
        return c


class FrameParams(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("max_bounce_count", C.c_uint32),
                ("samples_per_pixel", C.c_uint32), ("light_position", C.c_float * 3),
                ("light_intensity", C.c_float), ("output_format", C.c_uint32),
                ("output_location", C.c_uint32), ("shard_index", C.c_uint32), ("shard_count", C.c_uint32),
                ("tile_w", C.c_uint32), ("tile_h", C.c_uint32), ("shard_layout", C.c_uint32)]

    @staticmethod
    def make(width=800, height=600, max_bounce_count=63, samples_per_pixel=4, light_position=(5, 5, 5),
             light_intensity=1.0, output_format=FORMAT_RGBA8, output_location=LOCATION_HOST, shard_index=0,
             shard_count=1, tile_w=0, tile_h=0, shard_layout=SHARD_COMPACT) -> "FrameParams":
        p = FrameParams()
        p.width, p.height = int(width), int(height)
        p.max_bounce_count, p.samples_per_pixel = int(max_bounce_count), int(samples_per_pixel)
        p.light_position[:] = [float(x) for x in light_position]
        p.light_intensity = float(light_intensity)
        p.output_format, p.output_location = int(output_format), int(output_location)
        p.shard_index, p.shard_count = int(shard_index), int(shard_count)
        p.tile_w, p.tile_h = int(tile_w), int(tile_h)
        p.shard_layout = int(shard_layout)
        return p

    def copy(self) -> "FrameParams":
        q = FrameParams()
        C.memmove(C.byref(q), C.byref(self), C.sizeof(FrameParams))
        return q


class Material(C.Structure):
    """rt_material: Ka / Kd / Ks / Ns / Ni of one instance; the defaults are the constants of shader.rgen:51-55, :125."""
    _fields_ = [("ka", C.c_float * 3), ("kd", C.c_float * 3), ("ks", C.c_float * 3), ("ns", C.c_float), ("ni", C.c_float)]

    @staticmethod
    def make(ka=(0.1, 0.3, 0.1), kd=(0.2, 1.0, 0.2), ks=(0.8, 0.8, 0.8), ns=100.0, ni=1.52) -> "Material":
        m = Material()
        m.ka[:], m.kd[:], m.ks[:] = [float(x) for x in ka], [float(x) for x in kd], [float(x) for x in ks]
        m.ns, m.ni = float(ns), float(ni)
        return m


def material_array(materials):
    """list of Material -> contiguous ctypes array (what rt_scene_set_materials takes)."""
    arr = (Material * len(materials))()
    for i, m in enumerate(materials):
        C.memmove(C.byref(arr, i * C.sizeof(Material)), C.byref(m), C.sizeof(Material))
    return arr


class IpcHandle(C.Structure):
    _fields_ = [("opaque", C.c_uint8 * 64)]


class Hit(C.Structure):
    _fields_ = [("instance", C.c_uint32), ("primitive", C.c_uint32), ("t", C.c_float), ("u", C.c_float),
                ("v", C.c_float)]


HIT_DTYPE = np.dtype([("instance", np.uint32), ("primitive", np.uint32), ("t", np.float32), ("u", np.float32),
                      ("v", np.float32)])


class Stats(C.Structure):
    _fields_ = [("build_ms", C.c_double), ("frame_ms", C.c_double), ("trace_ms", C.c_double),
                ("shadow_ms", C.c_double), ("shade_ms", C.c_double),
                ("rays_primary", C.c_uint64), ("rays_secondary", C.c_uint64), ("rays_shadow", C.c_uint64),
                ("n_miss", C.c_uint64), ("n_hit", C.c_uint64), ("nodes_visited", C.c_uint64),
                ("tris_tested", C.c_uint64), ("kernel_launches", C.c_uint64), ("bvh_nodes", C.c_uint64),
                ("bvh_bytes", C.c_uint64), ("node_bytes", C.c_uint32), ("tri_bytes", C.c_uint32),
                ("ray_state_bytes", C.c_uint32), ("max_tree_depth", C.c_uint32), ("rays_top_level", C.c_uint64),
                ("stack_overflow", C.c_uint32), ("reserved", C.c_uint32)]

    def as_dict(self) -> dict:
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


@dataclasses.dataclass
class Config:
    """Runtime form of include/config.h:4-27 (+ window size main.cpp:805, light main.cpp:1853-1854)."""
    skybox_texture_dir: str = "resources/skybox_texture_sea"
    center_mesh_obj_path: str = "resources/teapot.obj"
    orbiting_mesh_obj_path: str = "resources/armadillo.obj"
    center_mesh_type: int = 1
    orbiting_mesh_type: int = 0
    camera_mouse_sensitivity: float = 0.0005
    camera_speed: float = 50.0
    max_bounce_count: int = 63
    samples_per_pixel: int = 4
    width: int = 800
    height: int = 600
    light_position: tuple = (5.0, 5.0, 5.0)
    light_intensity: float = 1.0


# ----------------------------------------------------------------------------- host lib
_host = None


def host_lib():
    global _host
    if _host is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise RtError(f"{HOST_LIB_PATH} not built; run __graft_entry__.build()")
        L = C.CDLL(HOST_LIB_PATH)
        fpp, upp, u32p = C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.POINTER(C.c_uint32)), C.POINTER(C.c_uint32)
        L.rth_last_error.restype = C.c_char_p
        L.rth_free.argtypes = [C.c_void_p]
        L.rth_load_obj.argtypes = [C.c_char_p, fpp, u32p, upp, u32p, C.POINTER(C.c_int)]
        L.rth_make_standin_armadillo.argtypes = [C.c_uint32, C.c_uint32, C.c_float, fpp, u32p, upp, u32p]
        L.rth_make_grid.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                    C.c_float, fpp, u32p, upp, u32p]
        L.rth_generate_normals.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]
        L.rth_load_mtl.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
        L.rth_load_skybox_dir.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_uint8)), u32p, u32p]
        L.rth_make_procedural_skybox.argtypes = [C.c_uint32, C.c_void_p]
        L.rth_camera_create.restype = C.c_void_p
        L.rth_camera_create.argtypes = [C.c_float] * 3
        L.rth_camera_destroy.argtypes = [C.c_void_p]
        L.rth_camera_move.argtypes = [C.c_void_p, C.c_int, C.c_float]
        L.rth_camera_mouse.argtypes = [C.c_void_p, C.c_float, C.c_float]
        L.rth_camera_look.argtypes = [C.c_void_p, C.c_int]
        L.rth_camera_get.argtypes = [C.c_void_p, C.c_void_p]
        L.rth_camera_view_matrix.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.rth_orbit_transforms.argtypes = [C.c_float, C.c_void_p]
        L.rth_write_ppm.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_void_p]
        _host = L
    return _host


def _hcheck(rc):
    if rc != 0:
        raise RtError(f"host library error {rc}: {host_lib().rth_last_error().decode()}")


def _take_mesh(v, nv, i, nt):
    L = host_lib()
    V = np.ctypeslib.as_array(v, (nv.value, 6)).copy() if nv.value else np.zeros((0, 6), np.float32)
    I = np.ctypeslib.as_array(i, (nt.value, 3)).copy() if nt.value else np.zeros((0, 3), np.uint32)
    L.rth_free(v)
    L.rth_free(i)
    return V, I


def load_obj(path: str):
    """OBJ -> (vertices[n,6] float32, indices[t,3] uint32, policy). src/main.cpp:1603-1729."""
    L = host_lib()
    v, i = C.POINTER(C.c_float)(), C.POINTER(C.c_uint32)()
    nv, nt, pol = C.c_uint32(), C.c_uint32(), C.c_int()
    _hcheck(L.rth_load_obj(path.encode(), C.byref(v), C.byref(nv), C.byref(i), C.byref(nt), C.byref(pol)))
    V, I = _take_mesh(v, nv, i, nt)
    return V, I, pol.value


def make_standin_armadillo(level: int = 7, seed: int = 20240611, radius: float = 2.5):
    L = host_lib()
    v, i = C.POINTER(C.c_float)(), C.POINTER(C.c_uint32)()
    nv, nt = C.c_uint32(), C.c_uint32()
    _hcheck(L.rth_make_standin_armadillo(level, seed, radius, C.byref(v), C.byref(nv), C.byref(i), C.byref(nt)))
    return _take_mesh(v, nv, i, nt)


def make_grid(V: np.ndarray, I: np.ndarray, nx: int, nz: int, pitch: float):
    L = host_lib()
    V = np.ascontiguousarray(V, np.float32)
    I = np.ascontiguousarray(I, np.uint32)
    v, i = C.POINTER(C.c_float)(), C.POINTER(C.c_uint32)()
    nv, nt = C.c_uint32(), C.c_uint32()
    _hcheck(L.rth_make_grid(V.ctypes.data, V.shape[0], I.ctypes.data, I.shape[0], nx, nz, pitch, C.byref(v),
                            C.byref(nv), C.byref(i), C.byref(nt)))
    return _take_mesh(v, nv, i, nt)


class HostMaterial(C.Structure):
    _fields_ = [("name", C.c_char * 64), ("ka", C.c_float * 3), ("kd", C.c_float * 3), ("ks", C.c_float * 3),
                ("ns", C.c_float), ("ni", C.c_float)]


def load_mtl(path: str):
    """Wavefront .mtl -> [(name, Material), ...] in file order (rth_load_mtl)."""
    L = host_lib()
    n = C.c_uint32(0)
    buf = (HostMaterial * 256)()
    _hcheck(L.rth_load_mtl(path.encode(), buf, 256, C.byref(n)))
    out = []
    for i in range(min(n.value, 256)):
        h = buf[i]
        out.append((h.name.decode(), Material.make(tuple(h.ka), tuple(h.kd), tuple(h.ks), h.ns, h.ni)))
    return out


def load_skybox_dir(path: str) -> np.ndarray:
    """6 JPEGs -> uint8[6,h,w,4], layer order right,left,top,bottom,front,back (main.cpp:2064-2071)."""
    L = host_lib()
    p, w, h = C.POINTER(C.c_uint8)(), C.c_uint32(), C.c_uint32()
    _hcheck(L.rth_load_skybox_dir(path.encode(), C.byref(p), C.byref(w), C.byref(h)))
    a = np.ctypeslib.as_array(p, (6, h.value, w.value, 4)).copy()
    L.rth_free(p)
    return a


def make_procedural_skybox(size: int = 2048) -> np.ndarray:
    a = np.zeros((6, size, size, 4), np.uint8)
    _hcheck(host_lib().rth_make_procedural_skybox(size, a.ctypes.data))
    return a


def orbit_transforms(time_param: float) -> np.ndarray:
    out = np.zeros((2, 12), np.float32)
    host_lib().rth_orbit_transforms(time_param, out.ctypes.data)
    return out


def write_ppm(path: str, rgba8: np.ndarray):
    a = np.ascontiguousarray(rgba8, np.uint8)
    _hcheck(host_lib().rth_write_ppm(path.encode(), a.shape[1], a.shape[0], a.ctypes.data))


RIGHT, LEFT, UP, DOWN, FORWARD, BACKWARD = range(6)


class Camera:
    """The reference's Camera (include/camera.h:16-35) through librt_host.so."""

    def __init__(self, position=(0.0, 0.0, 20.0)):
        self._h = host_lib().rth_camera_create(*[float(x) for x in position])

    def __del__(self):
        if getattr(self, "_h", None) and _host is not None:
            _host.rth_camera_destroy(self._h)
            self._h = None

    def _get(self):
        o = np.zeros(12, np.float32)
        host_lib().rth_camera_get(self._h, o.ctypes.data)
        return o.reshape(4, 3)

    def getPosition(self): return self._get()[0]
    def getRightVector(self): return self._get()[1]
    def getUpVector(self): return self._get()[2]
    def getFrontVector(self): return self._get()[3]
    def move(self, direction: int, distance: float): host_lib().rth_camera_move(self._h, direction, distance)
    def processMouseMovement(self, xoffset: float, yoffset: float): host_lib().rth_camera_mouse(self._h, xoffset, yoffset)
    def look(self, direction: int): host_lib().rth_camera_look(self._h, direction)

    def getViewingMatrix(self, without_translation=False):
        o = np.zeros(16, np.float32)
        host_lib().rth_camera_view_matrix(self._h, int(without_translation), o.ctypes.data)
        return o.reshape(4, 4).T  # column-major storage -> math layout

    def vectors(self) -> CameraVectors:
        g = self._get()
        return CameraVectors.make(g[0], g[1], g[2], g[3])


IDENTITY34 = (1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0)


def translation34(x, y, z):
    return (1, 0, 0, x, 0, 1, 0, y, 0, 0, 1, z)


class SceneInputs:
    """Host-side scene description that keeps the numpy buffers alive and produces an
    rt_scene_desc for either the product library or (in tests) the oracle."""

    def __init__(self, meshes: Sequence, instances: Sequence, skybox: Optional[np.ndarray] = None):
        self.meshes = [(np.ascontiguousarray(V, np.float32).reshape(-1, 6),
                        np.ascontiguousarray(I, np.uint32).reshape(-1, 3)) for V, I in meshes]
        self.instances = [(int(m), int(t), tuple(float(x) for x in xf)) for m, t, xf in instances]
        self.skybox = None if skybox is None else np.ascontiguousarray(skybox, np.uint8)
        if self.skybox is not None:
            assert self.skybox.ndim == 4 and self.skybox.shape[0] == 6 and self.skybox.shape[3] == 4

    def desc(self, device_id=-1, flags=0) -> SceneDesc:
        self._m = (MeshDesc * max(1, len(self.meshes)))()
        for k, (V, I) in enumerate(self.meshes):
            self._m[k].vertices = V.ctypes.data_as(C.POINTER(C.c_float))
            self._m[k].n_vertices = V.shape[0]
            self._m[k].indices = I.ctypes.data_as(C.POINTER(C.c_uint32))
            self._m[k].n_triangles = I.shape[0]
        self._i = (InstanceDesc * max(1, len(self.instances)))()
        for k, (m, t, xf) in enumerate(self.instances):
            self._i[k].mesh, self._i[k].type = m, t
            self._i[k].transform[:] = xf
        d = SceneDesc()
        d.n_meshes, d.meshes = len(self.meshes), self._m
        d.n_instances, d.instances = len(self.instances), self._i
        if self.skybox is not None:
            d.skybox_rgba8 = self.skybox.ctypes.data_as(C.POINTER(C.c_uint8))
            d.sky_h, d.sky_w = self.skybox.shape[1], self.skybox.shape[2]
        d.device_id, d.flags = device_id, flags
        return d


# ----------------------------------------------------------------------------- product lib
_rt = None


def rt_lib():
    """Loads librt_b200.so. Raises (never falls back) when it is missing."""
    global _rt
    if _rt is None:
        if not os.path.exists(RT_LIB_PATH):
            raise RtError(f"{RT_LIB_PATH} not built: the CUDA library is the only render path "
                          f"(run __graft_entry__.build())")
        L = C.CDLL(RT_LIB_PATH)
        L.rt_last_error.restype = C.c_char_p
        L.rt_abi_version.restype = C.c_uint32
        L.rt_scene_build.argtypes = [C.POINTER(SceneDesc), C.POINTER(C.c_void_p)]
        L.rt_scene_set_transforms.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.rt_scene_destroy.argtypes = [C.c_void_p]
        L.rt_generate_normals.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.POINTER(C.c_double)]
        L.rt_scene_set_materials.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.rt_scene_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.rt_set_counting.argtypes = [C.c_void_p, C.c_int]
        L.rt_render_frame.argtypes = [C.c_void_p, C.POINTER(CameraVectors), C.POINTER(FrameParams), C.c_void_p]
        L.rt_frame_output_bytes.restype = C.c_uint64
        L.rt_frame_output_bytes.argtypes = [C.POINTER(FrameParams)]
        L.rt_host_buffer_pin.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        L.rt_host_buffer_unpin.argtypes = [C.c_void_p]
        L.rt_frame_shard_pixels.restype = C.c_uint64
        L.rt_frame_shard_pixels.argtypes = [C.POINTER(FrameParams)]
        L.rt_assemble_shards.argtypes = [C.c_void_p, C.POINTER(FrameParams), C.c_void_p, C.c_void_p]
        L.rt_peer_frame_create.argtypes = [C.c_int, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(IpcHandle)]
        L.rt_peer_frame_open.argtypes = [C.c_int, C.POINTER(IpcHandle), C.POINTER(C.c_void_p)]
        L.rt_peer_frame_signal.argtypes = [C.c_int, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p]
        L.rt_peer_frame_wait.argtypes = [C.c_int, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.rt_peer_frame_close.argtypes = [C.c_int, C.c_void_p]
        L.rt_peer_frame_destroy.argtypes = [C.c_int, C.c_void_p]
        L.rt_host_frame_create.argtypes = [C.c_char_p, C.c_uint64, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.rt_host_frame_open.argtypes = [C.c_char_p, C.c_uint64, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.rt_host_frame_signal.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32]
        L.rt_host_frame_wait.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32]
        L.rt_host_frame_close.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int]
        L.rt_primary_hits.argtypes = [C.c_void_p, C.POINTER(CameraVectors), C.POINTER(FrameParams), C.c_void_p]
        L.rt_trace_rays.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_int,
                                    C.c_void_p]
        L.rt_get_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
        L.rt_measure_fp32_tflops.argtypes = [C.c_int, C.POINTER(C.c_double)]
        L.rt_debug_sort_pairs.argtypes = [C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rt_sort_pairs_noalloc.argtypes = [C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
        _rt = L
    return _rt


def _check(rc):
    if rc != 0:
        e = RtError(f"rt_b200 error {rc}: {rt_lib().rt_last_error().decode()}")
        e.code = rc
        raise e


def device_count() -> int:
    return int(rt_lib().rt_device_count())


def pixel_bytes(fmt: int) -> int:
    return 4 if fmt == FORMAT_RGBA8 else 16


class Scene:
    """A built scene on one B200: rt_scene_build … rt_scene_destroy."""

    def __init__(self, inputs: SceneInputs, device_id: int = -1, accel: int = ACCEL_DEFAULT, flags: int = 0):
        self.inputs = inputs
        self._h = C.c_void_p()
        d = inputs.desc(device_id, accel | flags)
        _check(rt_lib().rt_scene_build(C.byref(d), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            rt_lib().rt_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_transforms(self, xforms):
        a = np.ascontiguousarray(xforms, np.float32).reshape(-1, 12)
        _check(rt_lib().rt_scene_set_transforms(self._h, a.ctypes.data, a.shape[0]))

    def set_materials(self, materials):
        """One Material per instance, or None to return to the reference's shader constants."""
        if materials is None:
            _check(rt_lib().rt_scene_set_materials(self._h, None, 0))
        else:
            arr = material_array(materials)
            _check(rt_lib().rt_scene_set_materials(self._h, C.cast(arr, C.c_void_p), len(materials)))

    def set_stream(self, cuda_stream: int):
        _check(rt_lib().rt_scene_set_stream(self._h, C.c_void_p(cuda_stream)))

    def set_counting(self, on: bool):
        _check(rt_lib().rt_set_counting(self._h, int(on)))

    def render(self, camera: CameraVectors, params: FrameParams) -> np.ndarray:
        """Host-output frame: uint8[h,w,4] or float32[h,w,4] (full frame, shard_count<=1)."""
        p = params.copy()
        p.output_location = LOCATION_HOST
        dt = np.uint8 if p.output_format == FORMAT_RGBA8 else np.float32
        if p.shard_count > 1:
            out = np.zeros((int(rt_lib().rt_frame_shard_pixels(C.byref(p))), 4), dt)
        else:
            out = np.zeros((p.height, p.width, 4), dt)
        _check(rt_lib().rt_render_frame(self._h, C.byref(camera), C.byref(p), out.ctypes.data))
        return out

    def pin_host_buffer(self, array: np.ndarray):
        """rt_host_buffer_pin: later host-output frames rendered into `array` (render_into with its address and
        output_location = LOCATION_HOST) are stored by the resolve kernel directly, without a copy.  Keep the
        array alive until unpin_host_buffer() / close()."""
        self._pinned = array
        _check(rt_lib().rt_host_buffer_pin(self._h, array.ctypes.data, array.nbytes))

    def unpin_host_buffer(self):
        _check(rt_lib().rt_host_buffer_unpin(self._h))
        self._pinned = None

    def render_into(self, camera: CameraVectors, params: FrameParams, ptr: int):
        """Raw call: ptr is a host or device address per params.output_location."""
        _check(rt_lib().rt_render_frame(self._h, C.byref(camera), C.byref(params), C.c_void_p(ptr)))

    def assemble_shards(self, params: FrameParams, gathered_ptr: int, frame_ptr: int):
        _check(rt_lib().rt_assemble_shards(self._h, C.byref(params), C.c_void_p(gathered_ptr), C.c_void_p(frame_ptr)))

    def primary_hits(self, camera: CameraVectors, params: FrameParams) -> np.ndarray:
        out = np.zeros(params.width * params.height * params.samples_per_pixel, HIT_DTYPE)
        _check(rt_lib().rt_primary_hits(self._h, C.byref(camera), C.byref(params), out.ctypes.data))
        return out

    def trace_rays(self, origins, dirs, tmin=0.001, tmax=10000.0, any_hit=False) -> np.ndarray:
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        out = np.zeros(o.shape[0], HIT_DTYPE)
        _check(rt_lib().rt_trace_rays(self._h, o.shape[0], o.ctypes.data, d.ctypes.data, tmin, tmax, int(any_hit),
                                      out.ctypes.data))
        return out

    def stats(self) -> dict:
        s = Stats()
        _check(rt_lib().rt_get_stats(self._h, C.byref(s)))
        return s.as_dict()


class PeerFrame:
    """A frame buffer the other ranks of the box store their tiles into (rt_peer_frame_*).

    Owner: PeerFrame.create(device, nbytes) -> .handle_bytes() goes to the peers.
    Peer:  PeerFrame.open(device, handle_bytes).  `.ptr` is the device address to pass as the
    output of Scene.render_into with shard_layout=SHARD_IN_FRAME."""

    def __init__(self, device_id: int, ptr: int, nbytes: int, owner: bool, handle: "IpcHandle | None"):
        self.device_id, self.ptr, self.nbytes, self.owner, self._handle = device_id, ptr, nbytes, owner, handle

    @staticmethod
    def create(device_id: int, nbytes: int) -> "PeerFrame":
        d, h = C.c_void_p(), IpcHandle()
        _check(rt_lib().rt_peer_frame_create(device_id, nbytes, C.byref(d), C.byref(h)))
        return PeerFrame(device_id, d.value, nbytes, True, h)

    @staticmethod
    def open(device_id: int, handle_bytes: bytes, nbytes: int = 0) -> "PeerFrame":
        h = IpcHandle()
        C.memmove(C.byref(h), bytes(handle_bytes), 64)
        d = C.c_void_p()
        _check(rt_lib().rt_peer_frame_open(device_id, C.byref(h), C.byref(d)))
        return PeerFrame(device_id, d.value, nbytes, False, None)

    def handle_bytes(self) -> bytes:
        return bytes(self._handle.opaque)

    def signal(self, rank: int, value: int, stream: int = 0):
        """This rank's tiles of frame `value` are in place (ordered behind the render on `stream`)."""
        _check(rt_lib().rt_peer_frame_signal(self.device_id, C.c_void_p(self.ptr), self.nbytes, rank, value & 0xFFFFFFFF,
                                             C.c_void_p(stream)))

    def wait(self, n_ranks: int, value: int, stream: int = 0, timeout_ms: int = 2000):
        """Owner: block until all n_ranks flags have reached `value`."""
        _check(rt_lib().rt_peer_frame_wait(self.device_id, C.c_void_p(self.ptr), self.nbytes, n_ranks, value & 0xFFFFFFFF,
                                           timeout_ms, C.c_void_p(stream)))

    @property
    def __cuda_array_interface__(self):   # lets torch / cupy view the frame as uint8[nbytes]
        return {"shape": (self.nbytes,), "typestr": "|u1", "data": (self.ptr, False), "version": 2}

    def release(self):
        if self.ptr:
            L = rt_lib()
            _check((L.rt_peer_frame_destroy if self.owner else L.rt_peer_frame_close)(self.device_id, C.c_void_p(self.ptr)))
            self.ptr = 0


class HostFrame:
    """A frame in shared pinned host memory that every rank's k_resolve stores its tiles into
    directly (rt_host_frame_*): `.dev_ptr` is the output of Scene.render_into with
    shard_layout=SHARD_IN_FRAME, `.array(h, w)` the uint8 view of the pixels on the host."""

    def __init__(self, name: str, device_id: int, host_ptr: int, dev_ptr: int, nbytes: int, owner: bool):
        self.name, self.device_id, self.host_ptr, self.dev_ptr, self.nbytes, self.owner = name, device_id, host_ptr, dev_ptr, nbytes, owner

    @staticmethod
    def _map(fn, name: str, device_id: int, nbytes: int, owner: bool) -> "HostFrame":
        h, d = C.c_void_p(), C.c_void_p()
        _check(fn(name.encode(), nbytes, device_id, C.byref(h), C.byref(d)))
        return HostFrame(name, device_id, h.value, d.value, nbytes, owner)

    @staticmethod
    def create(name: str, device_id: int, nbytes: int) -> "HostFrame":
        return HostFrame._map(rt_lib().rt_host_frame_create, name, device_id, nbytes, True)

    @staticmethod
    def open(name: str, device_id: int, nbytes: int) -> "HostFrame":
        return HostFrame._map(rt_lib().rt_host_frame_open, name, device_id, nbytes, False)

    def array(self, *shape) -> np.ndarray:
        buf = (C.c_uint8 * self.nbytes).from_address(self.host_ptr)
        a = np.frombuffer(buf, dtype=np.uint8)
        return a.reshape(shape) if shape else a

    def signal(self, rank: int, value: int):
        _check(rt_lib().rt_host_frame_signal(C.c_void_p(self.host_ptr), self.nbytes, rank, value & 0xFFFFFFFF))

    def wait(self, n_ranks: int, value: int, timeout_ms: int = 2000):
        _check(rt_lib().rt_host_frame_wait(C.c_void_p(self.host_ptr), self.nbytes, n_ranks, value & 0xFFFFFFFF, timeout_ms))

    def release(self):
        if self.host_ptr:
            _check(rt_lib().rt_host_frame_close(self.name.encode(), C.c_void_p(self.host_ptr), self.nbytes, self.device_id,
                                                1 if self.owner else 0))
            self.host_ptr = self.dev_ptr = 0


def generate_normals(vertices: np.ndarray, indices: np.ndarray, device_id: int = 0):
    """Area-weighted vertex normals on the GPU (rt_generate_normals): returns (vertices with the normal
    slots filled, device milliseconds)."""
    v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 6).copy()
    i = np.ascontiguousarray(indices, np.uint32).reshape(-1, 3)
    ms = C.c_double(0.0)
    _check(rt_lib().rt_generate_normals(device_id, v.ctypes.data, v.shape[0], i.ctypes.data, i.shape[0], C.byref(ms)))
    return v, ms.value


def generate_normals_host(vertices: np.ndarray, indices: np.ndarray) -> np.ndarray:
    """The host loader's version of the same (rth_generate_normals)."""
    v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 6).copy()
    i = np.ascontiguousarray(indices, np.uint32).reshape(-1, 3)
    _hcheck(host_lib().rth_generate_normals(v.ctypes.data, v.shape[0], i.ctypes.data, i.shape[0]))
    return v


def measure_fp32_tflops(device_id: int = 0) -> float:
    out = C.c_double()
    _check(rt_lib().rt_measure_fp32_tflops(device_id, C.byref(out)))
    return out.value


def sort_pairs(keys: np.ndarray, vals: np.ndarray, device_id: int = 0):
    k = np.ascontiguousarray(keys, np.uint64)
    v = np.ascontiguousarray(vals, np.uint32)
    ko, vo = np.empty_like(k), np.empty_like(v)
    _check(rt_lib().rt_debug_sort_pairs(device_id, k.size, k.ctypes.data, v.ctypes.data, ko.ctypes.data, vo.ctypes.data))
    return ko, vo
