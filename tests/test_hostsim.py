"""Host simulation of the product's CRT_HD code (BVH build functors, traversal, shading) against the
oracle. Debugging aid for a container without a GPU; the GPU parity tests are in test_gpu_parity.py."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol

SEED = 1804289383


class HostSim:
    def __init__(self, scene_ptr, leaf_max=3):
        self.lib = ol.hostsim()
        self.h = self.lib.hs_scene_create(C.cast(scene_ptr, C.c_void_p), leaf_max)
        assert self.h
        info = np.zeros(4, np.uint32)
        self.lib.hs_scene_info(self.h, info.ctypes.data)
        self.num_nodes, self.num_prims, self.depth, self.rounds = [int(x) for x in info]
        self.lib.hs_scene_giants.restype = C.c_uint32
        self.lib.hs_scene_giants.argtypes = [C.c_void_p]
        self.num_giants = int(self.lib.hs_scene_giants(self.h))

    def nodes(self):
        p = self.lib.hs_scene_nodes(self.h)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), shape=(self.num_nodes, 20)).copy()

    def prims(self):
        p = self.lib.hs_scene_prims(self.h)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), shape=(self.num_prims, 12)).copy()

    def primary(self, cam, W, H, rng):
        ids = np.empty((H, W), np.int32)
        t = np.empty((H, W), np.float32)
        n = C.c_uint64()
        p = C.c_uint64()
        self.lib.hs_primary(self.h, C.cast(C.pointer(cam), C.c_void_p), W, H, rng.ctypes.data, ids.ctypes.data, t.ctypes.data,
                            C.cast(C.pointer(n), C.c_void_p), C.cast(C.pointer(p), C.c_void_p))
        return ids, t, n.value, p.value

    def render(self, cam, W, H, spp, bounces, rng):
        rad = np.empty((H, W, 3), np.float32)
        img = np.empty((H, W, 4), np.uint8)
        r, n, p = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self.lib.hs_render(self.h, C.cast(C.pointer(cam), C.c_void_p), W, H, spp, bounces, rng.ctypes.data, rad.ctypes.data,
                           img.ctypes.data, C.cast(C.pointer(r), C.c_void_p), C.cast(C.pointer(n), C.c_void_p),
                           C.cast(C.pointer(p), C.c_void_p))
        return rad, img, r.value

    def close(self):
        self.lib.hs_scene_destroy(self.h)


def check_wide_bvh(hs, num_prims_expected):
    """Structural invariants: every primitive in exactly one leaf, children counted once, 80-byte nodes."""
    nodes, prims = hs.nodes(), hs.prims()
    assert nodes.shape[1] * 4 == 80 and prims.shape[1] * 4 == 48
    assert hs.num_prims == num_prims_expected
    ids = prims[:, 9]
    assert sorted(ids.tolist()) == list(range(num_prims_expected))
    seen_nodes = np.zeros(hs.num_nodes, np.int32)
    seen_prims = np.zeros(hs.num_prims, np.int32)
    seen_nodes[0] = 1
    for k in range(hs.num_nodes):
        n = nodes[k]
        imask = n[3] >> 24
        child_base, prim_base = int(n[4]), int(n[5])
        meta = [(int(n[6 + (i // 4)]) >> (8 * (i % 4))) & 0xff for i in range(8)]
        rank = 0
        for i, m in enumerate(meta):
            if m == 0:
                assert not (imask >> i) & 1
                continue
            if (imask >> i) & 1:
                assert m == (1 << 5) | (24 + i)
                seen_nodes[child_base + rank] += 1
                rank += 1
            else:
                cnt = {1: 1, 3: 2, 7: 3}[m >> 5]
                off = m & 31
                assert off + cnt <= 24
                seen_prims[prim_base + off: prim_base + off + cnt] += 1
    assert (seen_nodes == 1).all()
    # giants (primitives that dwarf the rest of the scene) sit after the BVH's records, outside every leaf
    g = hs.num_giants
    assert 0 <= g <= 4 and (seen_prims[:hs.num_prims - g] == 1).all() and (seen_prims[hs.num_prims - g:] == 0).all()


@pytest.mark.parametrize("name,param,res,spp,bounces", [("cornell", 0, (64, 64), 2, 8), ("cornell_blob", 20, (64, 36), 2, 20),
                                                        ("spheres", 0, (48, 32), 2, 20), ("terrain", 24, (64, 36), 1, 20)])
def test_hostsim_is_bit_exact_with_oracle(crt, name, param, res, spp, bounces):
    sc = crt.builtin_scene(name, param)
    osc = ol.OracleScene(sc.pointer())
    hs = HostSim(sc.pointer())
    check_wide_bvh(hs, sc.num_triangles + sc.num_spheres)
    W, H = res
    rng = ol.rng_states(SEED, 0, W, H)
    oid, ot = osc.primary(sc.camera, W, H, SEED)
    hid, ht, nn, npr = hs.primary(sc.camera, W, H, rng)
    assert np.array_equal(oid, hid)
    assert np.array_equal(ot.view(np.uint32), ht.view(np.uint32))
    orad, oimg, orays = osc.render(sc.camera, W, H, spp, bounces, SEED)
    hrad, himg, hrays = hs.render(sc.camera, W, H, spp, bounces, rng.copy())
    assert orays == hrays
    assert np.array_equal(orad.view(np.uint32), hrad.view(np.uint32))
    assert np.array_equal(oimg, himg)
    hs.close()
    osc.close()
    sc.free()


def test_hostsim_ploc_rounds_of_both_kinds(crt):
    """8 768 triangles: the first PLOC rounds run one by one, the rounds from 2 048 clusters on in the single call of
    Exec::ploc_small (crt_bvh_driver.h; the product's k_ploc_small). The tree must hold every primitive exactly once and
    find the oracle's primary hits."""
    sc = crt.builtin_scene("cornell_blob", 48)
    assert sc.num_triangles > 4 * 2048
    hs = HostSim(sc.pointer())
    check_wide_bvh(hs, sc.num_triangles + sc.num_spheres)
    osc = ol.OracleScene(sc.pointer())
    rng = ol.rng_states(SEED, 0, 64, 36)
    oid, ot = osc.primary(sc.camera, 64, 36, SEED)
    hid, ht, _, _ = hs.primary(sc.camera, 64, 36, rng)
    assert np.array_equal(oid, hid) and np.array_equal(ot.view(np.uint32), ht.view(np.uint32))
    hs.close()
    osc.close()
    sc.free()


@pytest.mark.parametrize("leaf_max", [1, 2, 3])
def test_hostsim_leaf_sizes(crt, leaf_max):
    sc = crt.builtin_scene("cornell_blob", 12)
    hs = HostSim(sc.pointer(), leaf_max)
    check_wide_bvh(hs, sc.num_triangles + sc.num_spheres)
    osc = ol.OracleScene(sc.pointer())
    rng = ol.rng_states(SEED, 0, 40, 24)
    oid, ot = osc.primary(sc.camera, 40, 24, SEED)
    hid, ht, _, _ = hs.primary(sc.camera, 40, 24, rng)
    assert np.array_equal(oid, hid) and np.array_equal(ot.view(np.uint32), ht.view(np.uint32))
    hs.close()
    osc.close()
    sc.free()


def test_hostsim_degenerate_inputs(crt):
    # a single primitive, coincident primitives, a zero-area triangle
    mats = [(crt.MAT_LAMBERTIAN, (0.5, 0.5, 0.5), (0, 0, 0), 0.0, 1.0)]
    s, keep = crt.make_scene_struct([[0, 0, -1], [1, 0, -1], [0, 1, -1]], [[0, 1, 2]], [0], np.zeros((0, 4)), [], mats)
    hs = HostSim(C.pointer(s))
    check_wide_bvh(hs, 1)
    hs.close()
    pos = [[0, 0, -1], [1, 0, -1], [0, 1, -1], [0, 0, -1], [0, 0, -1]]
    idx = [[0, 1, 2]] * 5 + [[3, 3, 4]]
    s, keep = crt.make_scene_struct(pos, idx, [0] * 6, [[0, 0, -3, 1], [0, 0, -3, 1]], [0, 0], mats)
    hs = HostSim(C.pointer(s))
    check_wide_bvh(hs, 8)
    osc = ol.OracleScene(C.pointer(s))
    cam = crt.Camera()
    cam.position[:] = (0.2, 0.2, 1)
    cam.world_up[:] = (0, 1, 0)
    cam.yaw_deg, cam.pitch_deg, cam.vfov_deg, cam.aspect, cam.aperture, cam.focus_dist = -90, 0, 60, 1, 0, 1
    rng = ol.rng_states(SEED, 0, 32, 32)
    # mode=1: the reference's own BVH cannot see an axis-aligned coplanar leaf (see test_oracle.py)
    oid, ot = osc.primary(cam, 32, 32, SEED, mode=1)
    hid, ht, _, _ = hs.primary(cam, 32, 32, rng)
    assert (oid >= 0).any()
    # coincident primitives: equal t; the product's rule keeps the largest id, the reference keeps the
    # last one it visits, which for identical copies in input order is also the largest id
    assert np.array_equal(ot.view(np.uint32), ht.view(np.uint32))
    assert np.array_equal(oid, hid)
    hs.close()
    osc.close()
