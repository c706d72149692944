"""ctypes wrapper of oracle/liboracle.so (CPU restatement of the reference) and of the host
simulation tests/hostsim/libhostsim.so. TEST INFRASTRUCTURE: only tests/, smoke() and bench.py's
cpu_baseline / --impl reference legs import this."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
HOSTSIM_SO = os.path.join(ROOT, "tests", "hostsim", "libhostsim.so")


def _newer(src_list, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(s) and os.path.getmtime(s) > t for s in src_list)


def build_oracle():
    srcs = [os.path.join(ROOT, "oracle", f) for f in ("oracle.cpp", "oracle.h", "ref_math.h", "xorwow_host.h")]
    if _newer(srcs, ORACLE_SO):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], stdout=subprocess.DEVNULL)
    return ORACLE_SO


def build_hostsim():
    csrc = os.path.join(ROOT, "raytracer-cuda_b200", "csrc")
    srcs = [os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp")] + [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith(".h")]
    if _newer(srcs, HOSTSIM_SO):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-o", HOSTSIM_SO,
                               os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp")])
    return HOSTSIM_SO


_oracle = None


def oracle():
    global _oracle
    if _oracle is None:
        lib = C.CDLL(build_oracle())
        P = C.c_void_p
        lib.orc_scene_create.restype = P
        lib.orc_scene_create.argtypes = [P]
        lib.orc_scene_destroy.argtypes = [P]
        lib.orc_scene_permutation.argtypes = [P, P]
        lib.orc_render.restype = C.c_int
        lib.orc_render.argtypes = [P, P] + [C.c_int] * 8 + [C.c_uint64, C.c_uint64, P, C.c_int, C.c_int, C.c_int, P, P, P]
        lib.orc_primary.restype = C.c_int
        lib.orc_primary.argtypes = [P, P] + [C.c_int] * 6 + [C.c_uint64, C.c_uint64, C.c_int, C.c_int, P, P]
        lib.orc_xorwow.argtypes = [C.c_uint64, C.c_uint64, C.c_int, P, P, P]
        lib.orc_xorwow_matrix.argtypes = [C.c_int, P]
        lib.orc_camera_vectors.argtypes = [P, P]
        lib.orc_rng_states.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, P]
        _oracle = lib
    return _oracle


class OracleScene:
    def __init__(self, scene_ptr):
        """scene_ptr: ctypes pointer to a crt_scene (same struct layout as the product's)."""
        self.lib = oracle()
        self.h = self.lib.orc_scene_create(C.cast(scene_ptr, C.c_void_p))
        if not self.h:
            raise RuntimeError("orc_scene_create failed")

    def close(self):
        if self.h:
            self.lib.orc_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def permutation(self, num_triangles):
        perm = np.empty(num_triangles, np.int32)
        self.lib.orc_scene_permutation(self.h, perm.ctypes.data)
        return perm

    def render(self, cam, W, H, spp, max_bounces, seed, subseq_base=0, window=None, mode=0, threads=0, rng_state=None, fresh=True):
        x0, y0, w, h = window if window else (0, 0, W, H)
        rad = np.empty((h, w, 3), np.float32)
        img = np.empty((h, w, 4), np.uint8)
        rays = C.c_uint64(0)
        rs = rng_state.ctypes.data if rng_state is not None else None
        rc = self.lib.orc_render(self.h, C.cast(C.pointer(cam), C.c_void_p), W, H, x0, y0, w, h, spp, max_bounces, seed, subseq_base,
                                 rs, 1 if fresh else 0, mode, threads, rad.ctypes.data, img.ctypes.data, C.cast(C.pointer(rays), C.c_void_p))
        if rc != 0:
            raise RuntimeError("orc_render failed")
        return rad, img, rays.value

    def primary(self, cam, W, H, seed, subseq_base=0, window=None, mode=0, threads=0):
        x0, y0, w, h = window if window else (0, 0, W, H)
        ids = np.empty((h, w), np.int32)
        t = np.empty((h, w), np.float32)
        rc = self.lib.orc_primary(self.h, C.cast(C.pointer(cam), C.c_void_p), W, H, x0, y0, w, h, seed, subseq_base, mode, threads,
                                  ids.ctypes.data, t.ctypes.data)
        if rc != 0:
            raise RuntimeError("orc_primary failed")
        return ids, t


def xorwow(seed, subseq, n=16):
    st = np.empty(6, np.uint32)
    raw = np.empty(n, np.uint32)
    uni = np.empty(n, np.float32)
    oracle().orc_xorwow(seed, subseq, n, st.ctypes.data, raw.ctypes.data, uni.ctypes.data)
    return st, raw, uni


def rng_states(seed, base, W, H):
    """AoS (H, W, 6) uint32 states curand_init(seed, base+pixel, 0)."""
    out = np.empty((H, W, 6), np.uint32)
    oracle().orc_rng_states(seed, base, W, H, 0, out.ctypes.data)
    return out


def camera_vectors(cam):
    out = np.empty(19, np.float32)
    oracle().orc_camera_vectors(C.cast(C.pointer(cam), C.c_void_p), out.ctypes.data)
    return out


_hostsim = None


def hostsim():
    global _hostsim
    if _hostsim is None:
        lib = C.CDLL(build_hostsim())
        P = C.c_void_p
        lib.hs_scene_create.restype = P
        lib.hs_scene_create.argtypes = [P, C.c_int]
        lib.hs_scene_destroy.argtypes = [P]
        lib.hs_scene_info.argtypes = [P, P]
        lib.hs_scene_nodes.restype = P
        lib.hs_scene_nodes.argtypes = [P]
        lib.hs_scene_prims.restype = P
        lib.hs_scene_prims.argtypes = [P]
        lib.hs_primary.argtypes = [P, P, C.c_int, C.c_int, P, P, P, P, P]
        lib.hs_render.argtypes = [P, P, C.c_int, C.c_int, C.c_int, C.c_int, P, P, P, P, P, P]
        _hostsim = lib
    return _hostsim
