"""GPU parity: the CUDA path, called through the C-ABI, against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star): primary-hit ids bit-exact; matched-seed radiance within 1e-4 relative on
non-divergent paths (this implementation reproduces the reference's rounding sequence, so the tests ask for
bit-exact floats and only tolerate the few pixels where a double-precision libm call can differ); RGBA8 equal.
At full BASELINE sizes the oracle is checked on crops plus size-independent properties.
"""
import os

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
SEED = 1804289383
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def make_renderer(crt, sc, W, H, mode, aspect=None, leaf_max=0):
    r = crt.Renderer(W, H, mode=mode, bvh_leaf_max=leaf_max)
    r.set_scene(sc)
    cam = sc.camera
    if aspect is not None:
        cam.aspect = aspect
    r.update_camera(cam)
    r.initialize(SEED, 0)
    return r


def rel_err(a, b):
    return np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), 1e-3)


@pytest.mark.parametrize("name,param,W,H", [("cornell", 0, 256, 256), ("cornell_blob", 0, 320, 180), ("spheres", 0, 300, 200),
                                            ("terrain", 200, 320, 180)])
def test_primary_ids_bit_exact(crt, name, param, W, H):
    sc = crt.builtin_scene(name, param)
    osc = ol.OracleScene(sc.pointer())
    r = make_renderer(crt, sc, W, H, crt.MODE_MEGAKERNEL)
    ids, t = r.primary_ids(SEED, 0)
    oid, ot = osc.primary(sc.camera, W, H, SEED)
    assert np.array_equal(ids, oid), "%d differing pixels" % (ids != oid).sum()
    assert np.array_equal(t.view(np.uint32), ot.view(np.uint32))
    assert (ids >= 0).mean() > 0.5
    r.close()
    osc.close()
    sc.free()


@pytest.mark.parametrize("mode", ["mega", "wave"])
@pytest.mark.parametrize("name,param,W,H,spp,bounces", [("cornell", 0, 256, 256, 16, 8), ("cornell_blob", 0, 320, 180, 4, 20),
                                                        ("spheres", 0, 300, 200, 4, 20)])
def test_matched_seed_radiance(crt, mode, name, param, W, H, spp, bounces):
    sc = crt.builtin_scene(name, param)
    osc = ol.OracleScene(sc.pointer())
    r = make_renderer(crt, sc, W, H, crt.MODE_MEGAKERNEL if mode == "mega" else crt.MODE_WAVEFRONT)
    st = r.render(spp, bounces)
    acc = r.accum()
    img = r.image_rgba8()
    orad, oimg, orays = osc.render(sc.camera, W, H, spp, bounces, SEED)
    assert np.all(acc[..., 3] == spp)
    rad = acc[..., :3] * np.float32(1.0 / spp)
    exact = (rad.view(np.uint32) == orad.view(np.uint32)).all(axis=-1)
    # tolerance of the spec: 1e-4 relative on non-divergent paths. Divergent pixels (a discrete decision
    # flipped by a last-ulp difference in double pow) must be very rare.
    close = (rel_err(rad, orad) <= 1e-4).all(axis=-1)
    assert close.mean() > 0.9995, close.mean()
    assert exact.mean() > 0.999, exact.mean()
    assert (img != oimg).any(axis=-1).mean() < 5e-4
    assert abs(int(st.rays) - int(orays)) <= max(64, int(orays) // 10000)
    assert st.samples == W * H * spp
    r.close()
    osc.close()
    sc.free()


@pytest.mark.parametrize("W,H,spp", [(200, 120, 6), (256, 160, 24), (203, 117, 5)])
def test_wavefront_equals_megakernel_bitwise(crt, W, H, spp, monkeypatch):
    """Megakernel, wavefront with the bulk-copy-fed shade kernel (k_wf_shade_bulk: aligned 32-pixel runs through
    cp.async.bulk + mbarrier, everything else through its per-thread fallback) and wavefront with the per-thread shade
    kernel (CRT_SHADE_BULK=0) give the same bits. 203x117: W*H is not a multiple of 4, so the RNG planes are not
    16-byte aligned and every warp takes the fallback; 24 spp: pixels finish at different iterations, the queue gets
    holes and is compacted while other warps still run on aligned runs."""
    sc = crt.builtin_scene("cornell_blob", 48)
    out = []
    for mode, bulk in ((crt.MODE_MEGAKERNEL, None), (crt.MODE_WAVEFRONT, "1"), (crt.MODE_WAVEFRONT, "0")):
        if bulk is not None:
            monkeypatch.setenv("CRT_SHADE_BULK", bulk)
        r = make_renderer(crt, sc, W, H, mode, aspect=W / H)
        st = r.render(spp, 20)
        out.append((r.accum().copy(), r.image_rgba8().copy(), r.rng_state().copy(), int(st.rays)))
        r.close()
    for k in (1, 2):
        assert np.array_equal(out[0][0].view(np.uint32), out[k][0].view(np.uint32))
        assert np.array_equal(out[0][1], out[k][1])
        assert np.array_equal(out[0][2], out[k][2])
        assert out[0][3] == out[k][3]
    sc.free()


@pytest.mark.parametrize("name,param", [("cornell_blob", 0), ("spheres", 0), ("terrain", 300)])
def test_bvh_build_arena_and_per_allocation_paths_agree(crt, name, param, monkeypatch):
    """The build takes its temporaries from one arena (CudaExec::reserve_tmp) and runs the PLOC rounds on at most 2048
    clusters in one launch (k_ploc_small); CRT_BUILD_NO_ARENA=1 is the per-allocation path. Same tree (node count,
    depth), same primary ids and t bits either way."""
    sc = crt.builtin_scene(name, param)
    out = []
    for no_arena in (False, True):
        if no_arena:
            monkeypatch.setenv("CRT_BUILD_NO_ARENA", "1")
        r = make_renderer(crt, sc, 320, 180, crt.MODE_MEGAKERNEL, aspect=320 / 180)
        st = r.stats()
        ids, t = r.primary_ids(SEED, 0)
        out.append((int(st.bvh_nodes), int(st.bvh_depth), int(st.bvh_prims), ids.copy(), t.copy()))
        r.close()
    assert out[0][:3] == out[1][:3]
    assert np.array_equal(out[0][3], out[1][3]) and np.array_equal(out[0][4].view(np.uint32), out[1][4].view(np.uint32))
    assert (out[0][3] >= 0).mean() > 0.5
    sc.free()


@pytest.mark.parametrize("env", [{"CRT_WF_GRAPH": "1"}, {"CRT_WF_GRAPH": "1", "CRT_WF_PARTS": "1"},
                                 {"CRT_WF_GRAPH": "1", "CRT_WF_BLOCKED": "1", "CRT_WF_PARTS": "4"}, {"CRT_WF_GRAPH": "0", "CRT_WF_BLOCKED": "1", "CRT_WF_PARTS": "3"}])
def test_graph_replay_and_blocked_order_bitwise(crt, env, monkeypatch):
    """Batches of iterations replayed as CUDA graphs (captured the second time a batch signature is seen), and the
    part-major ("temporal blocking") launch order: same bits as direct launches, frame after frame (the graphs are
    reused by the second frame, whose kernel arguments are the same)."""
    sc = crt.builtin_scene("cornell_blob", 48)
    W, H = 256, 160
    monkeypatch.setenv("CRT_WF_GRAPH", "0")
    r = make_renderer(crt, sc, W, H, crt.MODE_WAVEFRONT, aspect=W / H)
    want = []
    for _ in range(2):
        st = r.render(12, 20)
        want.append((r.accum().copy(), r.rng_state().copy(), int(st.rays)))
        assert st.graph_launches == 0
    r.close()
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    monkeypatch.setenv("CRT_TAIL_LIVE", "0")     # batches with a part in its tail stages are launched directly: keep this small frame out of the tail
    r = make_renderer(crt, sc, W, H, crt.MODE_WAVEFRONT, aspect=W / H)
    for f in range(2):
        st = r.render(12, 20)
        assert np.array_equal(want[f][0].view(np.uint32), r.accum().view(np.uint32))
        assert np.array_equal(want[f][1], r.rng_state())
        assert want[f][2] == int(st.rays)
        if env["CRT_WF_GRAPH"] == "1":
            assert st.graph_launches > 0 and st.kernel_launches > 100
    r.close()
    sc.free()


@pytest.mark.parametrize("rounds,parts", [(1, 2), (3, 1), (2, 3)])
def test_parked_rays_resume_bitwise(crt, rounds, parts, monkeypatch):
    """extend parks the rays that are still in flight when its queue runs out (k_wf_extend "drain") and the next
    launch resumes them from the saved stack. Forced here on every launch of a small frame (a ray then needs
    many launches); the frame must equal the megakernel's bit for bit, ray count included."""
    sc = crt.builtin_scene("cornell_blob", 48)
    W, H = 200, 120
    r = make_renderer(crt, sc, W, H, crt.MODE_MEGAKERNEL)
    st = r.render(5, 20)
    want = (r.accum().copy(), r.image_rgba8().copy(), r.rng_state().copy(), int(st.rays))
    r.close()
    monkeypatch.setenv("CRT_DRAIN_ROUNDS", str(rounds))
    monkeypatch.setenv("CRT_DRAIN_MIN_LIVE", "0")
    monkeypatch.setenv("CRT_WF_PARTS", str(parts))
    monkeypatch.setenv("CRT_TAIL_LIVE", "0")             # no per-pixel tail kernel: every iteration is a wavefront one
    r = make_renderer(crt, sc, W, H, crt.MODE_WAVEFRONT)
    for _ in range(2):                                   # second frame: the pools are reused
        r.initialize(SEED, 0)
        st = r.render(5, 20)
        assert np.array_equal(want[0].view(np.uint32), r.accum().view(np.uint32))
        assert np.array_equal(want[1], r.image_rgba8())
        assert np.array_equal(want[2], r.rng_state())
        assert want[3] == int(st.rays)
    r.close()
    sc.free()


def test_parked_rays_full_size_default_settings(crt, monkeypatch):
    """Default settings park rays only on long queues and hand the last pixels of a part to k_wf_tail: a 2560x1440
    frame, against a run with parking and the tail kernel off, and one with only the tail kernel off."""
    sc = crt.builtin_scene("cornell_blob")
    W, H = 2560, 1440
    out = []
    for rounds, tail in (("0", "0"), (None, "0"), (None, None)):
        for key, val in (("CRT_DRAIN_ROUNDS", rounds), ("CRT_TAIL_LIVE", tail)):
            if val is None:
                monkeypatch.delenv(key, raising=False)
            else:
                monkeypatch.setenv(key, val)
        r = make_renderer(crt, sc, W, H, crt.MODE_WAVEFRONT)
        st = r.render(3, 20)
        out.append((r.accum().copy(), r.rng_state().copy(), int(st.rays), r.image_rgba8().copy()))
        r.close()
    for k in (1, 2):
        assert np.array_equal(out[0][0].view(np.uint32), out[k][0].view(np.uint32))
        assert np.array_equal(out[0][1], out[k][1])
        assert out[0][2] == out[k][2]
        assert np.array_equal(out[0][3], out[k][3])
    sc.free()


@pytest.mark.parametrize("tail_live,parts,rounds,iters,lanes_min", [(10**9, 2, None, 64, 1), (10**9, 1, 2, 1, 32), (6000, 2, 1, 3, 5),
                                                                    (2500, 3, None, 7, 1), (300, 1, None, 1000, 2)])
def test_tail_kernel_bitwise(crt, tail_live, parts, rounds, iters, lanes_min, monkeypatch):
    """k_wf_tail finishes the pixels of a part whose queue has become short, one thread per pixel, entering the
    megakernel's loop in the middle of a path (after a closing extend that also completes parked rays), in stages
    of a bounded number of path segments with a write-back and a queue compaction between them. Whatever
    the stage length, the lanes used per warp and the hand-over point (after the first batch; late, with a few hundred pixels left; with rays parked on every
    launch before it), the frame equals the megakernel's bit for bit: accumulation, image, RNG state, ray count."""
    sc = crt.builtin_scene("cornell_blob", 48)
    W, H = 200, 120
    r = make_renderer(crt, sc, W, H, crt.MODE_MEGAKERNEL)
    st = r.render(7, 20)
    want = (r.accum().copy(), r.image_rgba8().copy(), r.rng_state().copy(), int(st.rays))
    r.close()
    monkeypatch.setenv("CRT_TAIL_LIVE", str(tail_live))
    monkeypatch.setenv("CRT_TAIL_ITERS", str(iters))          # path segments per pixel and stage, then write-back + compaction
    monkeypatch.setenv("CRT_TAIL_LANES_MIN", str(lanes_min))  # lanes of a warp that take a pixel
    monkeypatch.setenv("CRT_WF_PARTS", str(parts))
    if rounds is not None:
        monkeypatch.setenv("CRT_DRAIN_ROUNDS", str(rounds))
        monkeypatch.setenv("CRT_DRAIN_MIN_LIVE", "0")
    r = make_renderer(crt, sc, W, H, crt.MODE_WAVEFRONT)
    for _ in range(2):
        r.initialize(SEED, 0)
        st = r.render(7, 20)
        assert np.array_equal(want[0].view(np.uint32), r.accum().view(np.uint32))
        assert np.array_equal(want[1], r.image_rgba8())
        assert np.array_equal(want[2], r.rng_state())
        assert want[3] == int(st.rays)
    # passes accumulate across the hand-over as well
    r.initialize(SEED, 0)
    r.render(3, 20)
    r.render(4, 20, flags=crt.RENDER_ACCUMULATE)
    assert np.array_equal(want[0][..., :3].view(np.uint32), r.accum()[..., :3].view(np.uint32))
    r.close()
    sc.free()


def test_rng_state_matches_curand_init(crt):
    sc = crt.builtin_scene("cornell")
    r = crt.Renderer(64, 48, mode=crt.MODE_MEGAKERNEL)
    r.initialize(SEED, 0)
    planes = r.rng_state()                       # (6, H, W)
    ref = ol.rng_states(SEED, 0, 64, 48)         # (H, W, 6)
    assert np.array_equal(np.moveaxis(planes, 0, -1), ref)
    r.initialize(SEED, 64 * 48 * 3)              # multi-GPU rank 3's disjoint subsequences
    ref = ol.rng_states(SEED, 64 * 48 * 3, 64, 48)
    assert np.array_equal(np.moveaxis(r.rng_state(), 0, -1), ref)
    r.close()
    sc.free()


@pytest.mark.parametrize("W,H,seed,base", [(2560, 1440, SEED, 0), (2560, 1440, SEED, 7 * 2560 * 1440), (203, 117, 0x0123456789abcdef, 12345),
                                          (64, 48, SEED, (1 << 40) + 77), (1, 1, 0, 0)])
def test_cooperative_rng_init_equals_curand_init(crt, W, H, seed, base, monkeypatch):
    """crt_init_rng computes the states cooperatively (k_rng_init_coop: one generic skip-ahead per CTA, then 255 matrix-vector
    products shared by its warps); CRT_RNG_INIT_CURAND=1 runs cuRAND's own curand_init per pixel. Every word of every pixel
    must agree: a full 2560x1440 frame from base 0 and from rank 7's base, an odd size with a 64-bit seed, a base beyond
    2^40 (digits 20 and up of the skip-ahead), a single pixel."""
    r = crt.Renderer(W, H, mode=crt.MODE_MEGAKERNEL)
    r.initialize(seed, base)
    coop = r.rng_state().copy()
    coop_ms = r.stats().rng_init_ms
    monkeypatch.setenv("CRT_RNG_INIT_CURAND", "1")
    r.initialize(seed, base)
    ref = r.rng_state()
    assert np.array_equal(coop, ref), "%d differing words" % (coop != ref).sum()
    assert len(np.unique(coop[4])) > min(W * H, 1000) // 2          # the states do differ from pixel to pixel
    print("rng init %dx%d: cooperative %.3f ms, curand_init %.3f ms" % (W, H, coop_ms, r.stats().rng_init_ms))
    r.close()


@pytest.mark.parametrize("mode", ["mega", "wave"])
def test_passes_accumulate_like_one_render(crt, mode):
    """spp split into passes (ACCUMULATE) continues the per-pixel streams: same bits as one call,
    like consecutive reference frames continue their RNG state (CUDAKernels.h:153,165)."""
    m = crt.MODE_MEGAKERNEL if mode == "mega" else crt.MODE_WAVEFRONT
    sc = crt.builtin_scene("cornell")
    a = make_renderer(crt, sc, 96, 96, m)
    a.render(12, 8)
    b = make_renderer(crt, sc, 96, 96, m)
    b.render(5, 8)
    b.render(4, 8, flags=crt.RENDER_ACCUMULATE)
    b.render(3, 8, flags=crt.RENDER_ACCUMULATE)
    assert np.array_equal(a.accum().view(np.uint32), b.accum().view(np.uint32))
    assert np.array_equal(a.image_rgba8(), b.image_rgba8())
    assert np.array_equal(a.rng_state(), b.rng_state())
    # without the flag a render overwrites (the reference always overwrites, CUDAKernels.h:164)
    b.render(2, 8)
    assert np.all(b.accum()[..., 3] == 2)
    a.close()
    b.close()
    sc.free()


def test_sample_sharding_sums_to_single_stream_statistics(crt):
    """Multi-GPU partition: rank g renders spp/G samples from subsequence g*W*H+pixel.
    Rank 0 alone must equal the single-GPU render of the same spp; the summed buffer is a valid
    estimate of the same image (checked against a longer single-stream render within MC error)."""
    sc = crt.builtin_scene("cornell")
    W = H = 64
    r = make_renderer(crt, sc, W, H, crt.MODE_WAVEFRONT)
    r.render(64, 8)
    ref = r.accum().copy()
    total = np.zeros_like(ref)
    for g in range(4):
        r.initialize(SEED, g * W * H)
        r.render(16, 8, flags=crt.RENDER_NO_TONEMAP)
        part = r.accum().copy()
        if g == 0:
            r0 = make_renderer(crt, sc, W, H, crt.MODE_WAVEFRONT)
            r0.render(16, 8)
            assert np.array_equal(part.view(np.uint32), r0.accum().view(np.uint32))
            r0.close()
        total += part
    assert np.all(total[..., 3] == 64)
    r.set_accum(total)
    r.tonemap()
    img = r.image_rgba8()
    assert img[..., 3].min() == 255
    m1 = ref[..., :3].mean() / 64
    m2 = total[..., :3].mean() / 64
    assert abs(m1 - m2) / m1 < 0.05
    r.close()
    sc.free()


def test_converged_image_within_monte_carlo_error(crt):
    """Converged images: per-pixel error within 1.5x the Monte-Carlo standard error. Two independent
    estimates (product with another seed vs the oracle's reference stream), each of spp samples:
    their difference has standard error sqrt(2) * sigma/sqrt(spp); sigma from the oracle's batches."""
    sc = crt.builtin_scene("cornell")
    W = H = 48
    spp, batches = 64, 8
    osc = ol.OracleScene(sc.pointer())
    rng = ol.rng_states(SEED, 0, W, H).reshape(-1)
    means = []
    for _ in range(batches):
        rad, _, _ = osc.render(sc.camera, W, H, spp, 8, SEED, rng_state=rng, fresh=False)
        means.append(rad.astype(np.float64))
    means = np.stack(means)
    omean = means.mean(axis=0)
    sem_batch = means.std(axis=0, ddof=1) / np.sqrt(batches)          # std error of the oracle's total mean
    r = make_renderer(crt, sc, W, H, crt.MODE_WAVEFRONT)
    r.initialize(777, 0)
    r.render(spp * batches, 8)
    acc = r.accum().astype(np.float64)
    pmean = acc[..., :3] / acc[..., 3:4]
    se_diff = np.sqrt(2.0) * sem_batch
    z = np.abs(pmean - omean) / np.maximum(se_diff, 1e-6)
    # emission is returned unweighted (fireflies), so per-pixel sigma is itself noisy: ask that the bulk of
    # pixels is inside 1.5 sigma-of-difference as a normal distribution predicts (86.6 %), with slack
    assert (z < 1.5).mean() > 0.80, (z < 1.5).mean()
    assert abs(pmean.mean() - omean.mean()) / omean.mean() < 0.02
    r.close()
    osc.close()
    sc.free()


def test_full_size_crops_and_properties(crt):
    """BASELINE config C2 at full 2560x1440: oracle on crops, plus size-independent properties."""
    sc = crt.builtin_scene("cornell_blob")
    W, H = 2560, 1440
    osc = ol.OracleScene(sc.pointer())
    r = make_renderer(crt, sc, W, H, crt.MODE_WAVEFRONT)
    ids, t = r.primary_ids(SEED, 0)
    for (x0, y0, w, h) in [(0, 0, 96, 48), (1200, 700, 128, 64), (2464, 1392, 96, 48), (700, 300, 64, 64)]:
        oid, ot = osc.primary(sc.camera, W, H, SEED, window=(x0, y0, w, h))
        assert np.array_equal(ids[y0:y0 + h, x0:x0 + w], oid)
        assert np.array_equal(t[y0:y0 + h, x0:x0 + w].view(np.uint32), ot.view(np.uint32))
    assert (ids >= 0).all()                       # the camera sits inside the closed box
    assert (ids >= 32).mean() > 0.01              # the blob is visible
    st = r.render(2, 20)
    acc = r.accum()
    assert np.all(acc[..., 3] == 2) and np.isfinite(acc).all() and (acc[..., :3] >= 0).all()
    x0, y0, w, h = 1100, 600, 96, 48
    orad, oimg, _ = osc.render(sc.camera, W, H, 2, 20, SEED, window=(x0, y0, w, h))
    rad = acc[y0:y0 + h, x0:x0 + w, :3] * np.float32(0.5)
    assert (rad.view(np.uint32) == orad.view(np.uint32)).all(axis=-1).mean() > 0.999
    assert st.rays > W * H * 2 * 2                # closed box: every sample bounces
    assert st.bvh_prims == sc.num_triangles + sc.num_spheres
    r.close()
    osc.close()
    sc.free()


def test_counted_render_reports_work(crt):
    sc = crt.builtin_scene("cornell_blob", 48)
    for mode in (crt.MODE_MEGAKERNEL, crt.MODE_WAVEFRONT):
        r = make_renderer(crt, sc, 160, 90, mode)
        a = r.render(2, 20, counted=True)
        assert a.nodes_fetched >= a.rays > 0 and a.prims_tested > 0
        rays_counted = int(a.rays)
        r.initialize(SEED, 0)
        b = r.render(2, 20)
        assert int(b.rays) == rays_counted and b.nodes_fetched == 0
        assert b.kernel_launches >= 1
        r.close()
    sc.free()


def test_error_paths_and_io(crt, tmp_path):
    sc = crt.builtin_scene("cornell")
    r = crt.Renderer(32, 32)
    with pytest.raises(crt.CrtError) as e:
        r.render(1, 8)
    assert e.value.code == crt.E_INVALID and "scene" in str(e.value)
    r.set_scene(sc)
    with pytest.raises(crt.CrtError):
        r.render(1, 8)                            # no camera
    r.update_camera(sc.camera)
    with pytest.raises(crt.CrtError):
        r.render(1, 8)                            # RNG not initialised
    r.initialize()
    with pytest.raises(crt.CrtError):
        r.render(0, 8)
    r.render(3, 8)
    ppm, pfm, ck = str(tmp_path / "a.ppm"), str(tmp_path / "a.pfm"), str(tmp_path / "a.ckpt")
    r.write_ppm(ppm)
    r.write_pfm(pfm)
    data = open(ppm, "rb").read()
    assert data.startswith(b"P6\n32 32\n255\n") and len(data) == len(b"P6\n32 32\n255\n") + 32 * 32 * 3
    img = r.image_rgba8()
    top_row = np.frombuffer(data[-32 * 32 * 3:], np.uint8).reshape(32, 32, 3)[0]
    assert np.array_equal(top_row, img[31, :, :3])           # file is top-down = flipVertically (WindowManager.h:89)
    # checkpoint / resume: save after 3 spp, continue 2 more; reload and continue -> same bits
    r.checkpoint_save(ck)
    r.render(2, 8, flags=crt.RENDER_ACCUMULATE)
    want = r.accum().copy()
    r2 = crt.Renderer(32, 32)
    r2.set_scene(sc)
    r2.update_camera(sc.camera)
    r2.checkpoint_load(ck)
    r2.render(2, 8, flags=crt.RENDER_ACCUMULATE)
    assert np.array_equal(want.view(np.uint32), r2.accum().view(np.uint32))
    r3 = crt.Renderer(16, 16)
    with pytest.raises(crt.CrtError):
        r3.checkpoint_load(ck)
    for x in (r, r2, r3):
        x.close()
    sc.free()


def test_empty_and_ragged_scenes(crt):
    mats = [(crt.MAT_LAMBERTIAN, (0.5, 0.5, 0.5), (0, 0, 0), 0.0, 1.0)]
    r = crt.Renderer(16, 16, mode=crt.MODE_MEGAKERNEL)
    s, keep = crt.make_scene_struct(np.zeros((0, 3)), np.zeros((0, 3)), [], np.zeros((0, 4)), [], mats)
    with pytest.raises(crt.CrtError) as e:
        r.set_scene(s)
    assert e.value.code == crt.E_INVALID
    s, keep = crt.make_scene_struct([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 7]], [0], np.zeros((0, 4)), [], mats)
    with pytest.raises(crt.CrtError):
        r.set_scene(s)                            # vertex index out of range
    # a negative or NaN albedo is refused (the wavefront kernels keep a flag in the sign bit of throughput.x); -0.0 is accepted
    for bad in ((-0.1, 0.5, 0.5), (0.5, float("nan"), 0.5), (0.5, 0.5, float("inf"))):
        s, keep = crt.make_scene_struct(np.zeros((0, 3)), np.zeros((0, 3)), [], [[0, 0, -3, 1]], [0], [(crt.MAT_LAMBERTIAN, bad, (0, 0, 0), 0.0, 1.0)])
        with pytest.raises(crt.CrtError) as e:
            r.set_scene(s)
        assert e.value.code == crt.E_INVALID and "albedo" in str(e.value)
    s, keep = crt.make_scene_struct(np.zeros((0, 3)), np.zeros((0, 3)), [], [[0, 0, -3, 1]], [0], [(crt.MAT_METAL, (-0.0, 0.5, 0.5), (0, 0, 0), 0.0, 1.0)])
    r.set_scene(s)
    # one sphere only, material index out of range: the reference neither scatters nor terminates
    # (CUDAKernels.h:127,135); the path burns its bounces and returns black
    s, keep = crt.make_scene_struct(np.zeros((0, 3)), np.zeros((0, 3)), [], [[0, 0, -3, 1]], [5], mats)
    r.set_scene(s)
    cam = crt.Camera()
    cam.position[:] = (0, 0, 0)
    cam.world_up[:] = (0, 1, 0)
    cam.yaw_deg, cam.pitch_deg, cam.vfov_deg, cam.aspect, cam.aperture, cam.focus_dist = -90, 0, 40, 1, 0, 1
    r.update_camera(cam)
    r.initialize()
    r.render(2, 8)
    acc = r.accum()
    assert np.all(acc[8, 8, :3] == 0) and acc[0, 0, :3].min() > 0.4     # centre black, corner sky
    import ctypes as C
    osc = ol.OracleScene(C.pointer(s))
    orad, _, _ = osc.render(cam, 16, 16, 2, 8, SEED)
    assert np.array_equal((acc[..., :3] * np.float32(0.5)).view(np.uint32), orad.view(np.uint32))
    osc.close()
    r.close()


def _small_camera(crt, pos=(0.0, 0.1, 2.5), vfov=40.0, aspect=1.0):
    cam = crt.Camera()
    cam.position[:] = pos
    cam.world_up[:] = (0, 1, 0)
    cam.yaw_deg, cam.pitch_deg, cam.vfov_deg, cam.aspect, cam.aperture, cam.focus_dist = -90, 0, vfov, aspect, 0.01, 2.5
    return cam


@pytest.mark.parametrize("W,H", [(37, 29), (1, 1), (130, 7)])
def test_odd_image_sizes_all_modes_equal_oracle(crt, W, H):
    """Sizes that are not multiples of the 16x16 tiles / of a warp; wavefront, megakernel and auto give the same bits."""
    sc = crt.builtin_scene("cornell")
    osc = ol.OracleScene(sc.pointer())
    cam = sc.camera
    cam.aspect = W / H
    orad, oimg, orays = osc.render(cam, W, H, 3, 8, SEED)
    for mode in (crt.MODE_WAVEFRONT, crt.MODE_MEGAKERNEL, crt.MODE_AUTO):
        r = crt.Renderer(W, H, mode=mode)
        r.set_scene(sc)
        r.update_camera(cam)
        r.initialize(SEED, 0)
        st = r.render(3, 8)
        rad = r.accum()[..., :3] / np.float32(3)
        assert np.array_equal(r.image_rgba8(), oimg)
        assert (np.abs(rad - orad) <= 1e-4 * np.maximum(np.abs(orad), 1e-3)).all()
        assert int(st.rays) == int(orays)
        r.close()
    osc.close()
    sc.free()


def test_giant_primitives_and_degenerate_triangles(crt):
    """Two huge spheres (ground + sky dome piece) next to a small mesh with a zero-area and a sliver triangle:
    the giants leave the BVH (stats.bvh_giants), results stay the oracle's bit for bit."""
    import ctypes as C
    rs = np.random.RandomState(5)
    pos = (rs.rand(60, 3).astype(np.float32) - 0.5) * 0.6
    idx = rs.randint(0, 60, size=(40, 3)).astype(np.uint32)
    idx[0] = (3, 3, 3)                 # zero-area triangle
    idx[1] = (4, 5, 4)                 # degenerate (two equal vertices)
    pos[7] = pos[8] + np.float32(1e-7)
    idx[2] = (7, 8, 9)                 # sliver
    tri_mat = rs.randint(0, 4, size=40).astype(np.int32)
    spheres = [[0, -1000.3, 0, 1000.0], [0, 0.1, 0, 0.12], [900.0, 900.0, -900.0, 700.0]]
    sph_mat = [0, 2, 1]
    mats = [(crt.MAT_LAMBERTIAN, (0.5, 0.5, 0.5), (0, 0, 0), 0.0, 1.0), (crt.MAT_METAL, (0.8, 0.6, 0.2), (0, 0, 0), 0.3, 1.0),
            (crt.MAT_DIELECTRIC, (1, 1, 1), (0, 0, 0), 0.0, 1.5), (crt.MAT_DIFFUSE_LIGHT, (0, 0, 0), (4, 4, 4), 0.0, 1.0)]
    s, keep = crt.make_scene_struct(pos, idx, tri_mat, spheres, sph_mat, mats, mesh_tri_offset=[0, 40])
    cam = _small_camera(crt)
    W, H = 96, 64
    cam.aspect = W / H
    osc = ol.OracleScene(C.pointer(s))
    orad, oimg, orays = osc.render(cam, W, H, 4, 12, SEED)
    oid, ot = osc.primary(cam, W, H, SEED)
    for mode in (crt.MODE_WAVEFRONT, crt.MODE_MEGAKERNEL):
        r = crt.Renderer(W, H, mode=mode)
        r.set_scene(s)
        assert r.stats().bvh_giants == 2 and r.stats().bvh_prims == 43
        r.update_camera(cam)
        r.initialize(SEED, 0)
        ids, t = r.primary_ids(SEED, 0)
        assert np.array_equal(ids, oid) and np.array_equal(t.view(np.uint32), ot.view(np.uint32))
        st = r.render(4, 12)
        rad = r.accum()[..., :3] * np.float32(0.25)
        exact = (rad.view(np.uint32) == orad.view(np.uint32)).all(axis=-1)
        assert exact.mean() > 0.995, exact.mean()
        assert (np.abs(rad - orad) <= 1e-4 * np.maximum(np.abs(orad), 1e-3)).all(axis=-1).mean() > 0.999
        assert abs(int(st.rays) - int(orays)) <= 16
        r.close()
    osc.close()


def test_second_set_scene_replaces_the_first(crt):
    a, b = crt.builtin_scene("cornell"), crt.builtin_scene("spheres")
    r = crt.Renderer(64, 48)
    r.set_scene(a)
    r.update_camera(a.camera)
    r.initialize(SEED, 0)
    r.render(2, 8)
    r.set_scene(b)
    r.update_camera(b.camera)
    r.initialize(SEED, 0)
    r.render(2, 20)
    osc = ol.OracleScene(b.pointer())
    orad, oimg, _ = osc.render(b.camera, 64, 48, 2, 20, SEED)
    assert (r.image_rgba8() != oimg).any(axis=-1).mean() < 1e-3
    assert r.stats().bvh_giants == 1             # the r=1000 ground sphere of the spheres scene
    osc.close()
    r.close()
    a.free()
    b.free()


def _write_obj_room(tmp_path):
    """A closed room with an emissive ceiling panel, a metal and a glass triangle pair: every material class of the
    reference's MTL classification (SceneManager.h:222-247), polygons that need fan triangulation, negative indices."""
    (tmp_path / "room.mtl").write_text(
        "newmtl white\nKd 0.73 0.73 0.73\nnewmtl red\nKd 0.65 0.05 0.05\nnewmtl lamp\nKe 15 15 15\nKd 0 0 0\n"
        "newmtl steel\nKd 0.8 0.8 0.9\nKs 0.5 0.5 0.5\nPr 0.1\nnewmtl glass\nKd 1 1 1\nd 0.3\nNi 1.5\n")
    v = ["v -1 -1 -1", "v 1 -1 -1", "v 1 1 -1", "v -1 1 -1", "v -1 -1 1", "v 1 -1 1", "v 1 1 1", "v -1 1 1",
         "v -0.3 0.99 -0.3", "v 0.3 0.99 -0.3", "v 0.3 0.99 0.3", "v -0.3 0.99 0.3",
         "v -0.6 -1 -0.2", "v -0.1 -1 -0.2", "v -0.35 -0.2 -0.4", "v 0.2 -1 0.3", "v 0.7 -1 0.3", "v 0.45 -0.3 0.1"]
    f = ["usemtl white", "f 1 2 6 5", "f 4 8 7 3", "f 1 4 3 2", "usemtl red", "f 1 5 8 4", "f 2 3 7 6",
         "usemtl lamp", "f 9 10 11 12", "usemtl steel", "f 13 14 15", "usemtl glass", "f -3 -2 -1"]
    (tmp_path / "room.obj").write_text("mtllib room.mtl\n" + "\n".join(v) + "\n" + "\n".join(f) + "\n")
    return str(tmp_path / "room.obj")


@pytest.mark.parametrize("mode", ["mega", "wave"])
def test_obj_scene_matches_oracle(crt, tmp_path, mode):
    """SURVEY.md 8(f) N1 through the render path: a scene that came in through crt_scene_load_obj (the reference's
    loadObject rules + its two hard-coded spheres) renders to the oracle's bits."""
    sc = crt.load_obj([_write_obj_room(tmp_path)])
    assert sc.num_triangles == 14 and sc.num_spheres == 2
    cam = crt.Camera()
    cam.position[:] = (0.0, 0.0, 0.55)
    cam.world_up[:] = (0.0, 1.0, 0.0)
    cam.yaw_deg, cam.pitch_deg, cam.vfov_deg, cam.aspect, cam.aperture, cam.focus_dist = -90.0, 0.0, 70.0, 1.5, 0.01, 0.5
    W, H, spp, bounces = 192, 128, 8, 12
    osc = ol.OracleScene(sc.pointer())
    r = crt.Renderer(W, H, mode=crt.MODE_MEGAKERNEL if mode == "mega" else crt.MODE_WAVEFRONT)
    r.set_scene(sc)
    r.update_camera(cam)
    r.initialize(SEED, 0)
    ids, t = r.primary_ids(SEED, 0)
    oid, ot = osc.primary(cam, W, H, SEED)
    assert np.array_equal(ids, oid)
    assert np.array_equal(t.view(np.uint32), ot.view(np.uint32))
    assert len(np.unique(ids)) >= 8          # walls, lamp, both triangles and the spheres are all in view
    st = r.render(spp, bounces)
    rad = r.accum()[..., :3] * np.float32(1.0 / spp)
    orad, oimg, orays = osc.render(cam, W, H, spp, bounces, SEED)
    assert int(st.rays) == int(orays)
    exact = (rad.view(np.uint32) == orad.view(np.uint32)).all(axis=-1)
    assert exact.mean() > 0.999, exact.mean()
    assert (rel_err(rad, orad) <= 1e-4).all(axis=-1).mean() > 0.9995
    assert (r.image_rgba8() != oimg).any(axis=-1).mean() < 5e-4
    assert rad.max() > 1.0                   # the lamp is seen directly
    r.close()
    osc.close()
    sc.free()


def test_async_render_and_wait(crt):
    """CRT_RENDER_ASYNC: crt_render returns at once, the frame runs on a host thread of the library; the
    context refuses other calls with CRT_E_BUSY until crt_wait, after which the frame is the synchronous one bit for bit."""
    sc = crt.builtin_scene("cornell_blob", 48)
    W, H = 640, 360
    r = make_renderer(crt, sc, W, H, crt.MODE_WAVEFRONT)
    st = r.render(16, 20)
    want = (r.accum().copy(), r.rng_state().copy(), int(st.rays))
    r.initialize(SEED, 0)
    assert r.render(16, 20, flags=crt.RENDER_ASYNC) is None
    with pytest.raises(crt.CrtError) as e:
        r.accum()
    assert e.value.code == crt.E_BUSY
    with pytest.raises(crt.CrtError):
        r.render(1, 8)
    st = r.wait()
    assert int(st.rays) == want[2]
    assert np.array_equal(r.accum().view(np.uint32), want[0].view(np.uint32))
    assert np.array_equal(r.rng_state(), want[1])
    assert r.wait().rays == st.rays          # nothing in flight: returns at once
    # a failing asynchronous render reports through crt_wait
    r2 = crt.Renderer(32, 32)
    r2.render(1, 8, flags=crt.RENDER_ASYNC)
    with pytest.raises(crt.CrtError) as e:
        r2.wait()
    assert e.value.code == crt.E_INVALID
    r2.close()
    r.close()
    sc.free()
