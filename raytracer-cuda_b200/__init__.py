"""ctypes binding of libcrt_b200.so — thin Python mirror of include/crt_api.h.

The product is the C-ABI shared library (CUDA kernels for sm_100a + C++ host code); this module
only loads it and marshals numpy buffers, for the tests and bench.py. It mirrors the reference's
operator surface for the render path (reference: CudaRayTracer/src/CUDARenderer.cuh:9-60:
initialize / updateCamera / render / getImageData / getRandState).

There is no Python or CPU fallback: if the library is missing or has no GPU, calls raise.
The directory name contains a hyphen, so import it with
    importlib.import_module("raytracer-cuda_b200")
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CRT_LIB") or os.path.join(_HERE, "libcrt_b200.so")   # CRT_LIB: tuning builds of the same library

CRT_OK = 0
E_INVALID, E_NO_DEVICE, E_CUDA, E_NOMEM, E_IO, E_PARSE, E_LIMIT, E_BUSY = -1, -2, -3, -4, -5, -6, -7, -8
MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_DIFFUSE_LIGHT = 0, 1, 2, 3
MODE_WAVEFRONT, MODE_MEGAKERNEL, MODE_AUTO = 0, 1, 2
RENDER_ACCUMULATE, RENDER_NO_TONEMAP, RENDER_TIME_KERNELS, RENDER_ASYNC = 1, 2, 4, 8
PRIMARY_STATIC_MAX_PIXELS = 1 << 20   # CRT_PRIMARY_STATIC_MAX_PIXELS
REFERENCE_SEED = 1804289383  # glibc rand() first value = what CUDARenderer::initialize passes (CUDARenderer.cuh:46)


class Material(C.Structure):
    _fields_ = [("type", C.c_int32), ("albedo", C.c_float * 3), ("emission", C.c_float * 3),
                ("roughness", C.c_float), ("ior", C.c_float)]


class Scene(C.Structure):
    _fields_ = [("positions", C.POINTER(C.c_float)), ("num_vertices", C.c_uint32),
                ("indices", C.POINTER(C.c_uint32)), ("tri_material", C.POINTER(C.c_int32)),
                ("num_triangles", C.c_uint32),
                ("spheres", C.POINTER(C.c_float)), ("sphere_material", C.POINTER(C.c_int32)),
                ("num_spheres", C.c_uint32),
                ("materials", C.POINTER(Material)), ("num_materials", C.c_uint32),
                ("mesh_tri_offset", C.POINTER(C.c_uint32)), ("num_meshes", C.c_uint32)]


class Camera(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("world_up", C.c_float * 3), ("yaw_deg", C.c_float),
                ("pitch_deg", C.c_float), ("vfov_deg", C.c_float), ("aspect", C.c_float),
                ("aperture", C.c_float), ("focus_dist", C.c_float)]


class Config(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("device", C.c_int32), ("mode", C.c_int32),
                ("bvh_leaf_max", C.c_int32), ("reference_exact", C.c_int32), ("reserved", C.c_int32 * 6)]


class RenderParams(C.Structure):
    _fields_ = [("spp", C.c_int32), ("max_bounces", C.c_int32), ("min_bounces", C.c_int32),
                ("rr_max_prob", C.c_float), ("flags", C.c_uint32), ("reserved", C.c_int32 * 3)]


class Stats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("rays_total", C.c_uint64), ("samples", C.c_uint64),
                ("render_ms", C.c_float), ("bvh_build_ms", C.c_float), ("rng_init_ms", C.c_float),
                ("scene_upload_ms", C.c_float), ("bvh_nodes", C.c_uint32), ("bvh_prims", C.c_uint32),
                ("bvh_depth", C.c_uint32), ("kernel_launches", C.c_uint32),
                ("nodes_fetched", C.c_uint64), ("prims_tested", C.c_uint64), ("extend_ms", C.c_float), ("shade_ms", C.c_float),
                ("extend_launches", C.c_uint32), ("shade_launches", C.c_uint32), ("bvh_giants", C.c_uint32),
                ("tail_ms", C.c_float), ("begin_ms", C.c_float), ("compact_ms", C.c_float), ("host_syncs", C.c_uint32),
                ("iterations", C.c_uint32), ("graph_launches", C.c_uint32), ("unseen_triangles", C.c_uint32)]


class CrtError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("crt error %d: %s" % (code, message))
        self.code = code


_lib = None


def load_library():
    """Load libcrt_b200.so (built by `make -C raytracer-cuda_b200`). Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError("%s not found: build it with __graft_entry__.build() / make -C raytracer-cuda_b200 "
                                "(there is no Python fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    P = C.c_void_p
    sig = {
        "crt_api_version": (C.c_int, []),
        "crt_create": (C.c_int, [C.POINTER(P), C.POINTER(Config)]),
        "crt_destroy": (None, [P]),
        "crt_last_error": (C.c_char_p, [P]),
        "crt_init_rng": (C.c_int, [P, C.c_uint64, C.c_uint64]),
        "crt_set_scene": (C.c_int, [P, C.POINTER(Scene)]),
        "crt_set_camera": (C.c_int, [P, C.POINTER(Camera)]),
        "crt_render": (C.c_int, [P, C.POINTER(RenderParams)]),
        "crt_render_counted": (C.c_int, [P, C.POINTER(RenderParams)]),
        "crt_wait": (C.c_int, [P]),
        "crt_primary_ids": (C.c_int, [P, C.c_uint64, C.c_uint64, P, P]),
        "crt_primary_ids_static": (C.c_int, [P, C.c_uint64, C.c_uint64, P, P]),
        "crt_get_image_rgba8": (C.c_int, [P, P, C.POINTER(P)]),
        "crt_get_accum": (C.c_int, [P, P, C.POINTER(P)]),
        "crt_set_accum": (C.c_int, [P, P]),
        "crt_tonemap": (C.c_int, [P]),
        "crt_get_rng_state": (C.c_int, [P, P, C.POINTER(P)]),
        "crt_set_rng_state": (C.c_int, [P, P]),
        "crt_get_stats": (C.c_int, [P, C.POINTER(Stats)]),
        "crt_get_stream": (C.c_int, [P, C.POINTER(P)]),
        "crt_set_stream": (C.c_int, [P, P]),
        "crt_scene_builtin": (C.c_int, [C.c_char_p, C.c_int32, C.POINTER(C.POINTER(Scene)), C.POINTER(Camera)]),
        "crt_scene_free": (None, [C.POINTER(Scene)]),
        "crt_scene_load_obj": (C.c_int, [C.POINTER(C.c_char_p), C.c_int32, C.c_int32, C.POINTER(C.POINTER(Scene)),
                                         C.c_char_p, C.c_size_t]),
        "crt_write_ppm": (C.c_int, [P, C.c_char_p]),
        "crt_write_pfm": (C.c_int, [P, C.c_char_p]),
        "crt_checkpoint_save": (C.c_int, [P, C.c_char_p]),
        "crt_checkpoint_load": (C.c_int, [P, C.c_char_p]),
        "crt_scene_reference_order": (C.c_int, [C.POINTER(Scene), P, P, P]),
        "crt_comm_unique_id": (C.c_int, [P, C.c_size_t]),
        "crt_comm_init": (C.c_int, [P, P, C.c_int32, C.c_int32]),
        "crt_comm_destroy": (C.c_int, [P]),
        "crt_comm_info": (C.c_int, [P, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
        "crt_reduce_accum": (C.c_int, [P, C.c_int32]),
        "crt_group_create": (C.c_int, [C.POINTER(P), C.POINTER(Config), C.POINTER(C.c_int32), C.c_int32]),
        "crt_group_destroy": (None, [P]),
        "crt_group_last_error": (C.c_char_p, [P]),
        "crt_group_size": (C.c_int32, [P]),
        "crt_group_ctx": (P, [P, C.c_int32]),
        "crt_group_set_scene": (C.c_int, [P, C.POINTER(Scene)]),
        "crt_group_set_camera": (C.c_int, [P, C.POINTER(Camera)]),
        "crt_group_init_rng": (C.c_int, [P, C.c_uint64]),
        "crt_group_render": (C.c_int, [P, C.POINTER(RenderParams), C.c_uint32]),
        "crt_group_get_stats": (C.c_int, [P, C.POINTER(Stats), C.POINTER(C.c_float), C.POINTER(C.c_int32)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)   # AttributeError here = the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


API_FUNCTIONS = ["crt_api_version", "crt_create", "crt_destroy", "crt_last_error", "crt_init_rng", "crt_set_scene",
                 "crt_set_camera", "crt_render", "crt_render_counted", "crt_wait", "crt_primary_ids", "crt_primary_ids_static", "crt_get_image_rgba8",
                 "crt_get_accum", "crt_set_accum", "crt_tonemap", "crt_get_rng_state", "crt_set_rng_state",
                 "crt_get_stats", "crt_get_stream", "crt_set_stream", "crt_scene_builtin", "crt_scene_free", "crt_scene_load_obj",
                 "crt_write_ppm", "crt_write_pfm", "crt_checkpoint_save", "crt_checkpoint_load",
                 "crt_scene_reference_order", "crt_comm_unique_id", "crt_comm_init", "crt_comm_destroy", "crt_comm_info", "crt_reduce_accum",
                 "crt_group_create", "crt_group_destroy", "crt_group_last_error", "crt_group_size", "crt_group_ctx",
                 "crt_group_set_scene", "crt_group_set_camera", "crt_group_init_rng", "crt_group_render", "crt_group_get_stats"]
COMM_ID_BYTES = 128
GROUP_FORCE_NCCL = 1


class SceneHandle:
    """A library-owned crt_scene (built-in or loaded from OBJ). Keeps numpy views of its arrays."""

    def __init__(self, ptr, camera=None):
        self._ptr = ptr
        self.camera = camera
        s = ptr.contents
        self.num_triangles, self.num_spheres = s.num_triangles, s.num_spheres
        self.num_vertices, self.num_materials, self.num_meshes = s.num_vertices, s.num_materials, s.num_meshes

    @property
    def struct(self):
        return self._ptr.contents

    def pointer(self):
        return self._ptr

    def arrays(self):
        s = self._ptr.contents
        out = {}
        out["positions"] = np.ctypeslib.as_array(s.positions, shape=(s.num_vertices, 3)).copy() if s.num_vertices else np.zeros((0, 3), np.float32)
        out["indices"] = np.ctypeslib.as_array(s.indices, shape=(s.num_triangles, 3)).copy() if s.num_triangles else np.zeros((0, 3), np.uint32)
        out["tri_material"] = np.ctypeslib.as_array(s.tri_material, shape=(s.num_triangles,)).copy() if s.num_triangles else np.zeros((0,), np.int32)
        out["spheres"] = np.ctypeslib.as_array(s.spheres, shape=(s.num_spheres, 4)).copy() if s.num_spheres else np.zeros((0, 4), np.float32)
        out["sphere_material"] = np.ctypeslib.as_array(s.sphere_material, shape=(s.num_spheres,)).copy() if s.num_spheres else np.zeros((0,), np.int32)
        out["mesh_tri_offset"] = np.ctypeslib.as_array(s.mesh_tri_offset, shape=(s.num_meshes + 1,)).copy() if s.num_meshes else np.zeros((0,), np.uint32)
        out["materials"] = [(m.type, tuple(m.albedo), tuple(m.emission), m.roughness, m.ior) for m in (s.materials[i] for i in range(s.num_materials))]
        return out

    def free(self):
        if self._ptr is not None:
            load_library().crt_scene_free(self._ptr)
            self._ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def builtin_scene(name, param=0):
    """crt_scene_builtin: 'cornell' (C1), 'cornell_blob' (C2/C5), 'spheres' (C3), 'terrain' (C4)."""
    lib = load_library()
    ptr = C.POINTER(Scene)()
    cam = Camera()
    rc = lib.crt_scene_builtin(name.encode(), int(param), C.byref(ptr), C.byref(cam))
    if rc != CRT_OK:
        raise CrtError(rc, "unknown built-in scene %r" % name)
    return SceneHandle(ptr, cam)


def load_obj(paths, add_reference_spheres=True):
    """crt_scene_load_obj: the reference's SceneManager::loadObject rules (SceneManager.h:198-329)."""
    lib = load_library()
    arr = (C.c_char_p * len(paths))(*[p.encode() for p in paths])
    ptr = C.POINTER(Scene)()
    err = C.create_string_buffer(512)
    rc = lib.crt_scene_load_obj(arr, len(paths), 1 if add_reference_spheres else 0, C.byref(ptr), err, 512)
    if rc != CRT_OK:
        raise CrtError(rc, err.value.decode(errors="replace"))
    return SceneHandle(ptr)


def make_scene_struct(positions, indices, tri_material, spheres, sphere_material, materials, mesh_tri_offset=None):
    """Build a crt_scene from numpy arrays. Returns (Scene, keepalive)."""
    pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
    idx = np.ascontiguousarray(indices, np.uint32).reshape(-1, 3)
    tm = np.ascontiguousarray(tri_material, np.int32).reshape(-1)
    sph = np.ascontiguousarray(spheres, np.float32).reshape(-1, 4)
    sm = np.ascontiguousarray(sphere_material, np.int32).reshape(-1)
    mats = (Material * max(len(materials), 1))()
    for i, (t, alb, emi, rough, ior) in enumerate(materials):
        mats[i].type = t
        mats[i].albedo[:] = alb
        mats[i].emission[:] = emi
        mats[i].roughness = rough
        mats[i].ior = ior
    s = Scene()
    s.positions = pos.ctypes.data_as(C.POINTER(C.c_float))
    s.num_vertices = pos.shape[0]
    s.indices = idx.ctypes.data_as(C.POINTER(C.c_uint32))
    s.tri_material = tm.ctypes.data_as(C.POINTER(C.c_int32))
    s.num_triangles = idx.shape[0]
    s.spheres = sph.ctypes.data_as(C.POINTER(C.c_float))
    s.sphere_material = sm.ctypes.data_as(C.POINTER(C.c_int32))
    s.num_spheres = sph.shape[0]
    s.materials = C.cast(mats, C.POINTER(Material))
    s.num_materials = len(materials)
    keep = [pos, idx, tm, sph, sm, mats]
    if mesh_tri_offset is not None and len(mesh_tri_offset) >= 2:
        off = np.ascontiguousarray(mesh_tri_offset, np.uint32)
        s.mesh_tri_offset = off.ctypes.data_as(C.POINTER(C.c_uint32))
        s.num_meshes = len(off) - 1
        keep.append(off)
    return s, keep


class Renderer:
    """CUDARenderer-shaped wrapper of a crt_ctx (reference: src/CUDARenderer.cuh:9-29)."""

    def __init__(self, width, height, device=0, mode=MODE_WAVEFRONT, bvh_leaf_max=0, borrowed_ctx=None, reference_exact=False):
        self.lib = load_library()
        self.width, self.height = int(width), int(height)
        self._borrowed = borrowed_ctx is not None
        if self._borrowed:                      # a context owned by a DeviceGroup
            self._ctx = C.c_void_p(borrowed_ctx)
            return
        cfg = Config(width=self.width, height=self.height, device=device, mode=mode, bvh_leaf_max=bvh_leaf_max,
                     reference_exact=1 if reference_exact else 0)
        self._ctx = C.c_void_p()
        rc = self.lib.crt_create(C.byref(self._ctx), C.byref(cfg))
        if rc != CRT_OK:
            msg = self.lib.crt_last_error(None).decode(errors="replace")
            self._ctx = None
            raise CrtError(rc, msg)

    def _check(self, rc):
        if rc == E_BUSY:
            raise CrtError(rc, "an asynchronous render owns the context: call wait()")
        if rc != CRT_OK:
            raise CrtError(rc, self.lib.crt_last_error(self._ctx).decode(errors="replace"))

    # CUDARenderer::initialize's initRandState launch
    def initialize(self, seed=REFERENCE_SEED, subsequence_base=0):
        self._check(self.lib.crt_init_rng(self._ctx, int(seed), int(subsequence_base)))

    def set_scene(self, scene):
        ptr = scene.pointer() if isinstance(scene, SceneHandle) else C.pointer(scene)
        self._check(self.lib.crt_set_scene(self._ctx, ptr))

    # CUDARenderer::updateCamera
    def update_camera(self, camera):
        self._check(self.lib.crt_set_camera(self._ctx, C.byref(camera)))

    # CUDARenderer::render
    def render(self, spp, max_bounces=20, min_bounces=3, rr_max_prob=0.95, flags=0, counted=False):
        p = RenderParams(spp=int(spp), max_bounces=int(max_bounces), min_bounces=int(min_bounces),
                         rr_max_prob=float(rr_max_prob), flags=int(flags))
        fn = self.lib.crt_render_counted if counted else self.lib.crt_render
        self._check(fn(self._ctx, C.byref(p)))
        return None if (int(flags) & RENDER_ASYNC) else self.stats()

    def wait(self):
        """crt_wait: join a render started with RENDER_ASYNC; returns its stats."""
        self._check(self.lib.crt_wait(self._ctx))
        return self.stats()

    def primary_ids(self, seed=REFERENCE_SEED, subsequence_base=0):
        ids = np.empty((self.height, self.width), np.int32)
        t = np.empty((self.height, self.width), np.float32)
        self._check(self.lib.crt_primary_ids(self._ctx, int(seed), int(subsequence_base), ids.ctypes.data, t.ctypes.data))
        return ids, t

    def primary_ids_static(self, seed=REFERENCE_SEED, subsequence_base=0, want_t=True):
        """crt_primary_ids_static: same buffers, staged in the library's static device array (no cudaMalloc /
        cudaFree inside the call); CrtError (E_LIMIT) above PRIMARY_STATIC_MAX_PIXELS."""
        ids = np.empty((self.height, self.width), np.int32)
        t = np.empty((self.height, self.width), np.float32) if want_t else None
        self._check(self.lib.crt_primary_ids_static(self._ctx, int(seed), int(subsequence_base), ids.ctypes.data,
                                                    t.ctypes.data if want_t else None))
        return ids, t

    # CUDARenderer::getImageData (+ the D2H copy of WindowManager::drawFrame); row 0 = bottom
    def image_rgba8(self, out=None):
        img = out if out is not None else np.empty((self.height, self.width, 4), np.uint8)
        self._check(self.lib.crt_get_image_rgba8(self._ctx, img.ctypes.data, None))
        return img

    def accum(self, out=None):
        acc = out if out is not None else np.empty((self.height, self.width, 4), np.float32)
        self._check(self.lib.crt_get_accum(self._ctx, acc.ctypes.data, None))
        return acc

    def accum_device_ptr(self):
        p = C.c_void_p()
        self._check(self.lib.crt_get_accum(self._ctx, None, C.byref(p)))
        return p.value

    def set_accum(self, acc):
        a = np.ascontiguousarray(acc, np.float32)
        assert a.size == self.width * self.height * 4
        self._check(self.lib.crt_set_accum(self._ctx, a.ctypes.data))

    def tonemap(self):
        self._check(self.lib.crt_tonemap(self._ctx))

    # CUDARenderer::getRandState as 6 planes (v0..v4, d)
    def rng_state(self):
        st = np.empty((6, self.height, self.width), np.uint32)
        self._check(self.lib.crt_get_rng_state(self._ctx, st.ctypes.data, None))
        return st

    def set_rng_state(self, st):
        a = np.ascontiguousarray(st, np.uint32)
        assert a.size == self.width * self.height * 6
        self._check(self.lib.crt_set_rng_state(self._ctx, a.ctypes.data))

    def stats(self):
        s = Stats()
        self._check(self.lib.crt_get_stats(self._ctx, C.byref(s)))
        return s

    def stream(self):
        p = C.c_void_p()
        self._check(self.lib.crt_get_stream(self._ctx, C.byref(p)))
        return p.value

    def set_stream(self, stream_ptr):
        self._check(self.lib.crt_set_stream(self._ctx, C.c_void_p(stream_ptr)))

    def write_ppm(self, path):
        self._check(self.lib.crt_write_ppm(self._ctx, path.encode()))

    def write_pfm(self, path):
        self._check(self.lib.crt_write_pfm(self._ctx, path.encode()))

    def checkpoint_save(self, path):
        self._check(self.lib.crt_checkpoint_save(self._ctx, path.encode()))

    def checkpoint_load(self, path):
        self._check(self.lib.crt_checkpoint_load(self._ctx, path.encode()))

    # ---- multi-GPU, one process per GPU ----
    def comm_init(self, unique_id, nranks, rank):
        """crt_comm_init: unique_id = the 128 bytes rank 0 got from comm_unique_id()."""
        buf = (C.c_char * COMM_ID_BYTES).from_buffer_copy(bytes(unique_id))
        self._check(self.lib.crt_comm_init(self._ctx, C.cast(buf, C.c_void_p), int(nranks), int(rank)))

    def comm_info(self):
        n, r, v = C.c_int32(), C.c_int32(), C.c_int32()
        self._check(self.lib.crt_comm_info(self._ctx, C.byref(n), C.byref(r), C.byref(v)))
        return n.value, r.value, v.value

    def reduce_accum(self, root=0):
        self._check(self.lib.crt_reduce_accum(self._ctx, int(root)))

    def comm_destroy(self):
        self._check(self.lib.crt_comm_destroy(self._ctx))

    def close(self):
        if self._ctx is not None and self._ctx.value and not self._borrowed:
            self.lib.crt_destroy(self._ctx)
        self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def reference_order(scene):
    """crt_scene_reference_order: (order, rank, unseen) of the reference builder's partition (host only, no GPU)."""
    lib = load_library()
    ptr = scene.pointer() if isinstance(scene, SceneHandle) else C.pointer(scene)
    T = ptr.contents.num_triangles
    order, rank, unseen = np.empty(T, np.uint32), np.empty(T, np.uint32), np.empty(T, np.uint8)
    rc = lib.crt_scene_reference_order(ptr, order.ctypes.data, rank.ctypes.data, unseen.ctypes.data)
    if rc != CRT_OK:
        raise CrtError(rc, "crt_scene_reference_order failed")
    return order, rank, unseen


def comm_unique_id():
    """crt_comm_unique_id (ncclGetUniqueId): 128 bytes for rank 0 to hand to the other ranks."""
    lib = load_library()
    buf = C.create_string_buffer(COMM_ID_BYTES)
    rc = lib.crt_comm_unique_id(C.cast(buf, C.c_void_p), COMM_ID_BYTES)
    if rc != CRT_OK:
        raise CrtError(rc, lib.crt_last_error(None).decode(errors="replace"))
    return buf.raw


class DeviceGroup:
    """crt_group_*: several GPUs driven from one process; device devices[0] ends up with the frame."""

    def __init__(self, width, height, devices, mode=MODE_WAVEFRONT, bvh_leaf_max=0):
        self.lib = load_library()
        self.width, self.height = int(width), int(height)
        cfg = Config(width=self.width, height=self.height, device=0, mode=mode, bvh_leaf_max=bvh_leaf_max)
        devs = (C.c_int32 * len(devices))(*[int(d) for d in devices])
        self._g = C.c_void_p()
        rc = self.lib.crt_group_create(C.byref(self._g), C.byref(cfg), devs, len(devices))
        if rc != CRT_OK:
            self._g = None
            raise CrtError(rc, self.lib.crt_last_error(None).decode(errors="replace"))

    def _check(self, rc):
        if rc != CRT_OK:
            raise CrtError(rc, self.lib.crt_group_last_error(self._g).decode(errors="replace"))

    def size(self):
        return self.lib.crt_group_size(self._g)

    def renderer(self, index):
        """The context of device `index` as a (borrowed) Renderer: read-backs, stats."""
        ctx = self.lib.crt_group_ctx(self._g, int(index))
        if not ctx:
            raise IndexError(index)
        return Renderer(self.width, self.height, borrowed_ctx=ctx)

    def set_scene(self, scene):
        ptr = scene.pointer() if isinstance(scene, SceneHandle) else C.pointer(scene)
        self._check(self.lib.crt_group_set_scene(self._g, ptr))

    def update_camera(self, camera):
        self._check(self.lib.crt_group_set_camera(self._g, C.byref(camera)))

    def initialize(self, seed=REFERENCE_SEED):
        self._check(self.lib.crt_group_init_rng(self._g, int(seed)))

    def render(self, spp, max_bounces=20, min_bounces=3, rr_max_prob=0.95, flags=0, group_flags=0):
        p = RenderParams(spp=int(spp), max_bounces=int(max_bounces), min_bounces=int(min_bounces),
                         rr_max_prob=float(rr_max_prob), flags=int(flags))
        self._check(self.lib.crt_group_render(self._g, C.byref(p), int(group_flags)))
        return self.stats()

    def stats(self):
        s, ms, peer = Stats(), C.c_float(), C.c_int32()
        self._check(self.lib.crt_group_get_stats(self._g, C.byref(s), C.byref(ms), C.byref(peer)))
        s.reduce_ms, s.peer = ms.value, bool(peer.value)
        return s

    def close(self):
        if self._g is not None and self._g.value:
            self.lib.crt_group_destroy(self._g)
        self._g = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
