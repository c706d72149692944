"""GPU parity at the FULL size of every BASELINE.json config (radiance, not only ids), and deliberate equal-t ties.

The oracle (CPU restatement of the reference, pinned by tests/golden) renders crops of the full-size frame with the
full-size camera; the CUDA path renders the whole frame through the C-ABI. A pixel's samples depend only on its own
cuRAND subsequence, so a crop of the full frame must carry the oracle's bits.
  C2 2560x1440 Cornell + 69 664-triangle blob   (reference: CUDAKernels.h:102-166, Core/Mesh.cuh:55-110)
  C3 1200x800  488 spheres, Metal fuzz / Dielectric (Core/Material.cuh:86-137, Core/BVHNode.cuh:115-156)
  C4 2560x1440 5 002 296-triangle terrain: the deep tree (826 k wide nodes), the 2*depth+2 <= 48 stack bound
  C5 1920x1080 1 spp, 8 bounces (the progressive frame)
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
SEED = 1804289383


def _renderer(crt, sc, W, H, mode):
    r = crt.Renderer(W, H, mode=mode)
    r.set_scene(sc)
    r.update_camera(sc.camera)
    r.initialize(SEED, 0)
    return r


def _check_crops(crt, sc, osc, W, H, spp, bounces, crops, mode, min_exact=0.999):
    r = _renderer(crt, sc, W, H, mode)
    ids, t = r.primary_ids(SEED, 0)
    st = r.render(spp, bounces)
    acc = r.accum()
    img = r.image_rgba8()
    assert np.all(acc[..., 3] == spp) and np.isfinite(acc).all()
    seen = set()
    for (x0, y0, w, h) in crops:
        oid, ot = osc.primary(sc.camera, W, H, SEED, window=(x0, y0, w, h), threads=8)
        assert np.array_equal(ids[y0:y0 + h, x0:x0 + w], oid), "primary ids differ in crop %r" % ((x0, y0, w, h),)
        assert np.array_equal(t[y0:y0 + h, x0:x0 + w].view(np.uint32), ot.view(np.uint32))
        seen.update(np.unique(oid).tolist())
        orad, oimg, _ = osc.render(sc.camera, W, H, spp, bounces, SEED, window=(x0, y0, w, h), threads=8)
        rad = acc[y0:y0 + h, x0:x0 + w, :3] * np.float32(1.0 / spp)
        exact = (rad.view(np.uint32) == orad.view(np.uint32)).all(axis=-1)
        # bit-exact but for the pixels where a last-ulp difference of double pow (Schlick) flips a decision
        assert exact.mean() >= min_exact, "crop %r: %.5f of the pixels bit-exact" % ((x0, y0, w, h), exact.mean())
        rel = np.abs(rad - orad) / np.maximum(np.maximum(np.abs(rad), np.abs(orad)), 1e-3)
        assert (rel <= 1e-4).all(axis=-1).mean() >= min_exact
        assert (img[y0:y0 + h, x0:x0 + w] != oimg).any(axis=-1).mean() <= 1.0 - min_exact
    r.close()
    return st, seen


@pytest.fixture(scope="module")
def terrain(crt):
    """The 5 002 296-triangle scene and the oracle's reference-order BVH over it (about half a minute of one core)."""
    sc = crt.builtin_scene("terrain", 0)
    osc = ol.OracleScene(sc.pointer())
    yield sc, osc
    osc.close()
    sc.free()


def test_c4_terrain_full_size_radiance(crt, terrain):
    """BASELINE configs[3] at its own size: 2560x1440 over 5 M triangles, 4 spp, 20 bounces, wavefront radiance against
    the oracle on crops that straddle the terrain silhouette, lie inside the terrain, and see only walls."""
    sc, osc = terrain
    assert sc.num_triangles == 5002296
    W, H = 2560, 1440
    crops = [(1408, 736, 96, 48),    # silhouette against the back wall, centre
             (128, 104, 96, 48),     # silhouette, left
             (1200, 200, 96, 48),    # inside the terrain
             (2464, 0, 96, 48),      # bottom-right corner
             (100, 1300, 64, 32)]    # walls only
    st, seen = _check_crops(crt, sc, osc, W, H, 4, 20, crops, crt.MODE_WAVEFRONT)
    assert st.bvh_prims == sc.num_triangles + sc.num_spheres
    assert 2 * st.bvh_depth + 2 <= 48                       # the traversal-stack bound crt_set_scene enforces
    assert st.bvh_nodes > 500000
    assert min(seen) < 64 and max(seen) > 1000000           # crops saw walls and the terrain
    assert len(seen) > 2000


def test_c4_terrain_megakernel_equals_wavefront(crt, terrain):
    sc, _ = terrain
    W, H = 640, 360
    out = []
    for mode in (crt.MODE_MEGAKERNEL, crt.MODE_WAVEFRONT):
        r = _renderer(crt, sc, W, H, mode)
        st = r.render(3, 20)
        out.append((r.accum().copy(), r.rng_state().copy(), int(st.rays)))
        r.close()
    assert np.array_equal(out[0][0].view(np.uint32), out[1][0].view(np.uint32))
    assert np.array_equal(out[0][1], out[1][1])
    assert out[0][2] == out[1][2]


@pytest.mark.parametrize("mode", ["mega", "wave"])
def test_c3_spheres_full_size_radiance(crt, mode):
    """BASELINE configs[2]: 1200x800, 20 bounces; 4 spp on crops over glass, metal and the ground sphere's horizon."""
    sc = crt.builtin_scene("spheres", 0)
    osc = ol.OracleScene(sc.pointer())
    crops = [(540, 360, 128, 64), (0, 0, 96, 48), (1104, 752, 96, 48), (300, 420, 96, 64), (800, 300, 96, 64)]
    st, seen = _check_crops(crt, sc, osc, 1200, 800, 4, 20, crops, crt.MODE_MEGAKERNEL if mode == "mega" else crt.MODE_WAVEFRONT,
                            min_exact=0.998)
    assert len(seen) > 8 and -1 in seen                     # several spheres and the sky
    osc.close()
    sc.free()


def test_c5_progressive_frame_full_size(crt):
    """BASELINE configs[4]: 1920x1080, 1 spp per frame, 8 bounces. Two consecutive frames continue the RNG streams
    (CUDAKernels.h:153,165); crops of both against the oracle."""
    sc = crt.builtin_scene("cornell_blob", 0)
    osc = ol.OracleScene(sc.pointer())
    W, H = 1920, 1080
    r = _renderer(crt, sc, W, H, crt.MODE_WAVEFRONT)
    crops = [(900, 500, 128, 64), (0, 0, 64, 32), (1856, 1048, 64, 32), (600, 250, 96, 64)]
    state = ol.rng_states(SEED, 0, W, H).reshape(-1)         # full-frame AoS states; the oracle continues them in place
    for frame in range(2):
        r.render(1, 8)
        acc = r.accum()
        assert np.all(acc[..., 3] == 1)
        for (x0, y0, w, h) in crops:
            orad, _, _ = osc.render(sc.camera, W, H, 1, 8, SEED, window=(x0, y0, w, h), rng_state=state, fresh=False, threads=8)
            rad = acc[y0:y0 + h, x0:x0 + w, :3]
            assert (rad.view(np.uint32) == orad.view(np.uint32)).all(axis=-1).mean() >= 0.999, (frame, x0, y0)
    # the device RNG state after two frames equals the oracle's on the crops
    planes = r.rng_state()                                   # (6, H, W)
    aos = state.reshape(H, W, 6)
    for (x0, y0, w, h) in crops:
        assert np.array_equal(np.moveaxis(planes[:, y0:y0 + h, x0:x0 + w], 0, -1), aos[y0:y0 + h, x0:x0 + w])
    r.close()
    osc.close()
    sc.free()


@pytest.mark.timeout(180, method="thread")
def test_consecutive_full_size_frames_finish(crt):
    """Regression test for the tail stall: in one code generation k_wf_tail never returned from the
    first tail stage of the 3rd or 4th consecutive headline-size frame. Five 1000-spp frames of BASELINE configs[1]'s
    scene at 2560x1440 (each frame hands both pixel parts to the tail kernel once, with the other part's extend / shade
    still running); every frame must finish (pytest-timeout ends the process otherwise), cover every pixel and trace a
    plausible number of rays."""
    sc = crt.builtin_scene("cornell_blob", 0)
    W, H, spp = 2560, 1440, 1000
    r = _renderer(crt, sc, W, H, crt.MODE_WAVEFRONT)
    rays = []
    for frame in range(5):
        st = r.render(spp, 20)
        rays.append(int(st.rays))
        assert st.kernel_launches > 1000
    acc = r.accum()
    assert np.all(acc[..., 3] == spp)
    assert np.isfinite(acc).all()
    mean = np.mean(rays)
    assert all(abs(x - mean) < 1e-4 * mean for x in rays), rays      # the frames continue the RNG streams: same count to 0.01 %
    assert 3.0 < mean / (W * H * spp) < 3.6                            # 3.26 path segments per sample on this scene
    r.close()
    sc.free()


# ---- equal-t ties -----------------------------------------------------------------------------------------------
def _tie_scene(crt, same_material):
    """A floor quad, and above it a fan of triangles each present TWICE (same three vertices, so the two copies give
    bit-identical t on every ray). The copies' ids are interleaved and reversed so that id order, input order and the
    reference builder's permuted order all differ. Copy B carries another material unless same_material."""
    rs = np.random.RandomState(11)
    pos = [[-1, 0, -1], [1, 0, -1], [1, 0, 1], [-1, 0, 1]]
    idx = [[0, 1, 2], [0, 2, 3]]
    mat = [0, 0]
    base = []
    for k in range(24):
        c = (rs.rand(3) - 0.5) * np.array([1.4, 0.0, 1.4]) + np.array([0, 0.25 + 0.02 * k, 0])
        tri = [c + (rs.rand(3) - 0.5) * 1.1 for _ in range(3)]
        v0 = len(pos)
        pos.extend([list(p) for p in tri])
        base.append([v0, v0 + 1, v0 + 2])
    for k, tri in enumerate(base):          # copy A
        idx.append(tri)
        mat.append(1 + (k % 2))
    for k, tri in reversed(list(enumerate(base))):   # copy B, reversed order
        idx.append(tri)
        mat.append(1 + (k % 2) if same_material else 3)
    mats = [(crt.MAT_LAMBERTIAN, (0.7, 0.7, 0.7), (0, 0, 0), 0.0, 1.0), (crt.MAT_LAMBERTIAN, (0.7, 0.2, 0.2), (0, 0, 0), 0.0, 1.0),
            (crt.MAT_METAL, (0.8, 0.8, 0.3), (0, 0, 0), 0.2, 1.0), (crt.MAT_DIFFUSE_LIGHT, (0, 0, 0), (3, 3, 3), 0.0, 1.0)]
    s, keep = crt.make_scene_struct(np.array(pos, np.float32), np.array(idx, np.uint32), mat, [[0, -100.5, 0, 100.0]], [0], mats,
                                    mesh_tri_offset=[0, len(idx)])
    cam = crt.Camera()
    cam.position[:] = (0.0, 2.2, 2.4)
    cam.world_up[:] = (0, 1, 0)
    cam.yaw_deg, cam.pitch_deg, cam.vfov_deg, cam.aspect, cam.aperture, cam.focus_dist = -90, -40, 50, 1.5, 0.0, 3.0
    twin = {}
    n0 = 2
    for k in range(24):
        a, b = n0 + k, n0 + 24 + (23 - k)
        twin[a], twin[b] = b, a
    return s, keep, cam, twin


@pytest.mark.parametrize("mode", ["mega", "wave"])
def test_equal_t_ties_duplicate_triangles(crt, mode):
    """Deliberate exact ties. The reference keeps the LATER equal-t hit in its own traversal order (Mesh.cuh:73-74,293:
    left-first over its permuted index array, inclusive interval); this library cannot follow that order through a
    different BVH and keeps the LARGER primitive id (crt_trace.h test_prim; DESIGN.md "Known reference quirks").
    What must hold: t bit-exact everywhere; ids equal wherever there is no tie; on a tie pixel the two ids are the two
    copies of one triangle and the product's is the larger; with equal materials on both copies the radiance is the
    oracle's bit for bit in spite of the id difference."""
    W, H = 192, 128
    m = crt.MODE_MEGAKERNEL if mode == "mega" else crt.MODE_WAVEFRONT
    s, keep, cam, twin = _tie_scene(crt, same_material=True)
    osc = ol.OracleScene(C.pointer(s))
    r = crt.Renderer(W, H, mode=m)
    r.set_scene(s)
    r.update_camera(cam)
    r.initialize(SEED, 0)
    ids, t = r.primary_ids(SEED, 0)
    oid, ot = osc.primary(cam, W, H, SEED)
    assert np.array_equal(t.view(np.uint32), ot.view(np.uint32))
    tie = np.isin(oid, list(twin.keys()))
    assert tie.mean() > 0.05                                  # the duplicated fan is well in view
    assert np.array_equal(ids[~tie], oid[~tie])
    pid, rid = ids[tie], oid[tie]
    partner = np.vectorize(twin.get)(rid)
    assert np.all((pid == rid) | (pid == partner))
    assert np.all(pid == np.maximum(rid, partner))            # the documented rule: larger id among equal t
    # how often the reference's own order picks the larger id too (for the record; printed with -s)
    print("tie pixels: %d, reference also keeps the larger id on %.1f %%" % (tie.sum(), 100.0 * (pid == rid).mean()))
    st = r.render(4, 12)
    rad = r.accum()[..., :3] * np.float32(0.25)
    orad, oimg, orays = osc.render(cam, W, H, 4, 12, SEED)
    assert np.array_equal(rad.view(np.uint32), orad.view(np.uint32))
    assert int(st.rays) == int(orays)
    r.close()
    osc.close()
    # different materials on the two copies: the divergence becomes visible exactly on the tie pixels whose winner
    # differs, nowhere else at sample 0 (4 spp: a path may meet a tie at any bounce, so only statistics are asked)
    s, keep, cam, twin = _tie_scene(crt, same_material=False)
    osc = ol.OracleScene(C.pointer(s))
    r = crt.Renderer(W, H, mode=m)
    r.set_scene(s)
    r.update_camera(cam)
    r.initialize(SEED, 0)
    ids, t = r.primary_ids(SEED, 0)
    oid, ot = osc.primary(cam, W, H, SEED)
    assert np.array_equal(t.view(np.uint32), ot.view(np.uint32))
    tie = np.isin(oid, list(twin.keys()))
    assert np.array_equal(ids[~tie], oid[~tie])
    r.close()
    osc.close()


def test_shared_edges_and_vertices_of_a_mesh(crt):
    """The realistic tie: rays through shared edges / vertices of a closed mesh. A pinhole camera at a lattice point
    looks at a regular grid whose vertices project exactly onto pixel centres is not possible with the reference's
    jittered camera, so the mesh is made fine instead (a 96x96 grid seen at 128x96: ~1 vertex per pixel) and every
    pixel is compared: ids must equal the oracle's wherever the two neighbours' t differ, and t always."""
    n = 96
    xs = np.linspace(-1, 1, n + 1, dtype=np.float32)
    gx, gz = np.meshgrid(xs, xs)
    gy = (0.05 * np.sin(7 * gx) * np.cos(5 * gz)).astype(np.float32)
    pos = np.stack([gx, gy, gz], -1).reshape(-1, 3)
    idx = []
    for j in range(n):
        for i in range(n):
            a = j * (n + 1) + i
            idx.append([a, a + 1, a + n + 2])
            idx.append([a, a + n + 2, a + n + 1])
    rs = np.random.RandomState(3)
    idx = np.array(idx, np.uint32)[rs.permutation(len(idx))]      # id order unrelated to position
    mats = [(crt.MAT_LAMBERTIAN, (0.6, 0.6, 0.6), (0, 0, 0), 0.0, 1.0)]
    s, keep = crt.make_scene_struct(pos, idx, [0] * len(idx), np.zeros((0, 4)), [], mats, mesh_tri_offset=[0, len(idx)])
    cam = crt.Camera()
    cam.position[:] = (0.0, 1.6, 1.9)
    cam.world_up[:] = (0, 1, 0)
    cam.yaw_deg, cam.pitch_deg, cam.vfov_deg, cam.aspect, cam.aperture, cam.focus_dist = -90, -40, 55, 4 / 3, 0.0, 2.0
    W, H = 128, 96
    osc = ol.OracleScene(C.pointer(s))
    r = crt.Renderer(W, H, mode=crt.MODE_WAVEFRONT)
    r.set_scene(s)
    r.update_camera(cam)
    r.initialize(SEED, 0)
    ids, t = r.primary_ids(SEED, 0)
    oid, ot = osc.primary(cam, W, H, SEED)
    assert np.array_equal(t.view(np.uint32), ot.view(np.uint32))
    assert (ids != oid).mean() < 1e-3              # a differing id can only be an exact edge tie
    assert (oid >= 0).mean() > 0.3
    st = r.render(8, 6)
    rad = r.accum()[..., :3] * np.float32(0.125)
    orad, _, orays = osc.render(cam, W, H, 8, 6, SEED)
    assert (rad.view(np.uint32) == orad.view(np.uint32)).all(axis=-1).mean() > 0.999
    r.close()
    osc.close()


# ---- reference-exact mode (crt_config.reference_exact) ---------------------------------------------------------------
@pytest.mark.parametrize("mode", ["mega", "wave"])
def test_reference_exact_ties_follow_the_reference_order(crt, mode):
    """With reference_exact the tie key is the triangle's position in the reference builder's permuted order
    (crt_scene_reference_order): on the duplicate-triangle scene, where the default rule disagrees with the reference on
    ~30 % of the tie pixels, ids equal the oracle's on EVERY pixel and, with different materials on the two copies,
    radiance is the oracle's bit for bit."""
    W, H = 192, 128
    m = crt.MODE_MEGAKERNEL if mode == "mega" else crt.MODE_WAVEFRONT
    s, keep, cam, twin = _tie_scene(crt, same_material=False)
    osc = ol.OracleScene(C.pointer(s))
    r = crt.Renderer(W, H, mode=m, reference_exact=True)
    r.set_scene(s)
    r.update_camera(cam)
    r.initialize(SEED, 0)
    ids, t = r.primary_ids(SEED, 0)
    oid, ot = osc.primary(cam, W, H, SEED)
    assert np.array_equal(t.view(np.uint32), ot.view(np.uint32))
    assert np.array_equal(ids, oid)
    st = r.render(4, 12)
    rad = r.accum()[..., :3] * np.float32(0.25)
    orad, oimg, orays = osc.render(cam, W, H, 4, 12, SEED)
    assert np.array_equal(rad.view(np.uint32), orad.view(np.uint32))
    assert int(st.rays) == int(orays) and st.unseen_triangles == 0
    r.close()
    # the default rule on the same scene: a different winner on part of the tie pixels, hence other radiance there
    r = crt.Renderer(W, H, mode=m)
    r.set_scene(s)
    r.update_camera(cam)
    r.initialize(SEED, 0)
    ids0, _ = r.primary_ids(SEED, 0)
    assert (ids0 != oid).mean() > 0.005
    r.close()
    osc.close()


@pytest.mark.parametrize("mode", ["mega", "wave"])
def test_reference_exact_hides_what_the_reference_never_sees(crt, mode):
    """A room of axis-aligned quads, each a mesh of its own: the reference's zero-thickness leaf boxes hide all of them
    (Mesh.cuh:258-264, AABB.cuh:135-136). reference_exact reproduces that: ids and radiance equal the faithful oracle;
    the default mode renders the room (stated divergence, DESIGN.md "Known reference quirks")."""
    from test_refexact import quad_room
    W, H = 192, 128
    m = crt.MODE_MEGAKERNEL if mode == "mega" else crt.MODE_WAVEFRONT
    s, keep, cam = quad_room(crt)
    osc = ol.OracleScene(C.pointer(s))
    oid, ot = osc.primary(cam, W, H, SEED)                     # faithful traversal
    bid, bt = osc.primary(cam, W, H, SEED, mode=1)             # brute force: what is really there
    r = crt.Renderer(W, H, mode=m, reference_exact=True)
    r.set_scene(s)
    assert r.stats().unseen_triangles == 12
    r.update_camera(cam)
    r.initialize(SEED, 0)
    ids, t = r.primary_ids(SEED, 0)
    assert np.array_equal(ids, oid) and np.array_equal(t.view(np.uint32), ot.view(np.uint32))
    st = r.render(4, 10)
    rad = r.accum()[..., :3] * np.float32(0.25)
    orad, _, orays = osc.render(cam, W, H, 4, 10, SEED)
    assert (rad.view(np.uint32) == orad.view(np.uint32)).all(axis=-1).mean() > 0.999
    assert int(st.rays) == int(orays)
    r.close()
    r = crt.Renderer(W, H, mode=m)                             # default: every triangle is seen
    r.set_scene(s)
    r.update_camera(cam)
    r.initialize(SEED, 0)
    ids0, t0 = r.primary_ids(SEED, 0)
    assert np.array_equal(ids0, bid) and np.array_equal(t0.view(np.uint32), bt.view(np.uint32))
    assert (ids0 != oid).mean() > 0.3 and r.stats().unseen_triangles == 0
    r.close()
    osc.close()
