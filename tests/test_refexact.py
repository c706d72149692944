"""Reference-exact mode, host side (no GPU): crt_scene_reference_order replays the PARTITION of the reference's mesh BVH
builder (CudaRayTracer/src/Core/Mesh.cuh:121-219). Pinned against the builder permutations dumped from the unmodified
reference kernels on a B200 (tests/golden/*_perm_i32.bin) and against the oracle's restatement of the same builder."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as ol

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SEED = 1804289383


@pytest.mark.parametrize("prefix,name,param", [("c1", "cornell", 0), ("c2s", "cornell_blob", 0)])
def test_order_equals_the_reference_kernels_permutation(crt, prefix, name, param):
    perm = np.fromfile(os.path.join(GOLDEN, prefix + "_perm_i32.bin"), dtype=np.int32)
    sc = crt.builtin_scene(name, param)
    order, rank, unseen = crt.reference_order(sc)
    assert np.array_equal(order.astype(np.int32), perm)
    assert np.array_equal(rank[order], np.arange(sc.num_triangles, dtype=np.uint32))
    assert unseen.sum() == 0            # the benchmark scenes lose no triangle to the reference's flat leaves
    sc.free()


@pytest.mark.parametrize("name,param", [("terrain", 120), ("cornell_blob", 31), ("terrain", 333)])
def test_order_equals_oracle_builder(crt, name, param):
    sc = crt.builtin_scene(name, param)
    osc = ol.OracleScene(sc.pointer())
    order, _, _ = crt.reference_order(sc)
    assert np.array_equal(order.astype(np.int32), osc.permutation(sc.num_triangles))
    osc.close()
    sc.free()


def quad_room(crt, per_quad_meshes=True):
    """A Cornell-like room of axis-aligned quads (2 triangles each), every quad its own mesh the way an OBJ with one
    shape per wall arrives (SceneManager.h:100-196), plus a tilted quad and a box made of 12 triangles in one mesh."""
    pos, idx, mat, off = [], [], [], [0]

    def quad(p0, p1, p2, p3, m, close=True):
        b = len(pos)
        pos.extend([p0, p1, p2, p3])
        idx.extend([[b, b + 1, b + 2], [b, b + 2, b + 3]])
        mat.extend([m, m])
        if close and per_quad_meshes:
            off.append(len(idx))

    quad([-1, -1, -1], [1, -1, -1], [1, -1, 1], [-1, -1, 1], 0)      # floor   (y = -1): flat leaf
    quad([-1, 1, -1], [-1, 1, 1], [1, 1, 1], [1, 1, -1], 0)          # ceiling (y = +1): flat leaf
    quad([-1, -1, -1], [-1, 1, -1], [1, 1, -1], [1, -1, -1], 0)      # back    (z = -1): flat leaf
    quad([-1, -1, -1], [-1, -1, 1], [-1, 1, 1], [-1, 1, -1], 1)      # left    (x = -1): flat leaf
    quad([1, -1, -1], [1, 1, -1], [1, 1, 1], [1, -1, 1], 2)          # right   (x = +1): flat leaf
    quad([-0.3, 0.99, -0.3], [0.3, 0.99, -0.3], [0.3, 0.99, 0.3], [-0.3, 0.99, 0.3], 3)   # lamp (y = .99): flat leaf
    quad([-0.6, -0.6, -0.2], [0.0, -0.5, -0.2], [0.0, 0.1, -0.5], [-0.6, 0.0, -0.5], 1)   # tilted: seen
    c, h = np.array([0.45, -0.7, 0.1]), 0.3                            # a box: 6 quads in ONE mesh -> one leaf of 12, seen
    for ax in range(3):
        for sgn in (-1, 1):
            u, v = [(1, 2), (0, 2), (0, 1)][ax]
            p = []
            for du, dv in ((-1, -1), (1, -1), (1, 1), (-1, 1)):
                q = c.copy(); q[ax] += sgn * h; q[u] += du * h; q[v] += dv * h
                p.append(list(q))
            quad(*p, 0, close=False)
    if per_quad_meshes:
        off.append(len(idx))
    mats = [(crt.MAT_LAMBERTIAN, (0.73, 0.73, 0.73), (0, 0, 0), 0.0, 1.0), (crt.MAT_LAMBERTIAN, (0.65, 0.05, 0.05), (0, 0, 0), 0.0, 1.0),
            (crt.MAT_LAMBERTIAN, (0.12, 0.45, 0.15), (0, 0, 0), 0.0, 1.0), (crt.MAT_DIFFUSE_LIGHT, (0, 0, 0), (15, 15, 15), 0.0, 1.0)]
    s, keep = crt.make_scene_struct(np.array(pos, np.float32), np.array(idx, np.uint32), mat, [[0.0, -0.75, -0.4, 0.25]], [0], mats,
                                    mesh_tri_offset=off if per_quad_meshes else [0, len(idx)])
    cam = crt.Camera()
    cam.position[:] = (0.0, 0.0, 2.6)
    cam.world_up[:] = (0, 1, 0)
    cam.yaw_deg, cam.pitch_deg, cam.vfov_deg, cam.aspect, cam.aperture, cam.focus_dist = -90, 0, 45, 1.5, 0.0, 2.6
    return s, keep, cam


def test_unseen_triangles_are_the_ones_the_oracle_never_hits(crt):
    """Flat leaves: the six axis-aligned quads that are meshes of their own. The faithful oracle traversal (mode 0) never
    reports them, brute force over all triangles (mode 1) does; the mask says exactly that."""
    s, keep, cam = quad_room(crt)
    order, rank, unseen = crt.reference_order(s)
    assert unseen.tolist() == [1] * 12 + [0] * 14
    osc = ol.OracleScene(C.pointer(s))
    W, H = 120, 80
    faithful, _ = osc.primary(cam, W, H, SEED, mode=0)
    brute, _ = osc.primary(cam, W, H, SEED, mode=1)
    hidden = set(np.flatnonzero(unseen).tolist())
    assert not (set(np.unique(faithful).tolist()) & hidden)
    assert set(np.unique(brute).tolist()) & hidden
    # where brute force sees a visible triangle first, so does the reference
    vis = ~np.isin(brute, list(hidden))
    assert np.array_equal(faithful[vis], brute[vis])
    osc.close()
    # the same room as ONE mesh: leaves mix planes, nothing is hidden
    s1, keep1, _ = quad_room(crt, per_quad_meshes=False)
    _, _, unseen1 = crt.reference_order(s1)
    assert unseen1.sum() == 0


def test_bad_mesh_offsets_are_rejected(crt):
    s, keep, _ = quad_room(crt)
    off = np.array([0, 4, 3, s.num_triangles], np.uint32)
    s.mesh_tri_offset = off.ctypes.data_as(C.POINTER(C.c_uint32))
    s.num_meshes = 3
    with pytest.raises(crt.CrtError):
        crt.reference_order(s)
