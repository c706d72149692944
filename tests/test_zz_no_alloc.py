"""crt_primary_ids_static: the allocation-free variant of crt_primary_ids (staging in a file-scope device array of the
library, csrc/crt_render.cu) must return exactly what crt_primary_ids returns. Collected last on purpose
(test_zz_*): a failure here must not hide the other tests.

Runs alone in a fresh process as well as after the other tests: it builds its own contexts and depends on no
state left behind (the static array keeps the last call's contents; the second scene below is smaller than the
first, so stale ids of the first frame sit right behind the second frame's pixels)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
SEED = 1804289383


def _renderer(crt, sc, W, H):
    r = crt.Renderer(W, H, mode=crt.MODE_MEGAKERNEL)
    r.set_scene(sc)
    cam = sc.camera
    cam.aspect = W / H
    r.update_camera(cam)
    return r


def _same(a, b):
    ids_a, t_a = a
    ids_b, t_b = b
    assert np.array_equal(ids_a, ids_b), "%d differing ids" % (ids_a != ids_b).sum()
    assert np.array_equal(t_a.view(np.uint32), t_b.view(np.uint32))   # bit-exact, the bar of every id / t test here


def test_static_scratch_entry_point_matches_crt_primary_ids(crt):
    # call 1: Cornell + blob, 320x180; call 2 straight after it: another scene, another (smaller, odd) size, another
    # subsequence base, so that a result that merely survived in the shared array cannot pass for the second call
    sc1 = crt.builtin_scene("cornell_blob", 0)
    r1 = _renderer(crt, sc1, 320, 180)
    ref1 = r1.primary_ids(SEED, 0)
    got1 = r1.primary_ids_static(SEED, 0)
    _same(got1, ref1)
    assert (ref1[0] >= 0).mean() > 0.5

    sc2 = crt.builtin_scene("spheres", 0)
    r2 = _renderer(crt, sc2, 301, 97)
    got2 = r2.primary_ids_static(SEED, 12345)
    ref2 = r2.primary_ids(SEED, 12345)
    _same(got2, ref2)
    assert not np.array_equal(got2[0], got1[0][:97, :301])

    # the first context again (two calls in a row on it), ids only (out_t = NULL)
    ids_only, none_t = r1.primary_ids_static(SEED, 0, want_t=False)
    assert none_t is None and np.array_equal(ids_only, ref1[0])
    _same(r1.primary_ids_static(SEED, 0), ref1)
    for r in (r1, r2):
        r.close()
    sc1.free()
    sc2.free()


def test_static_scratch_entry_point_rejects_frames_beyond_its_maximum(crt):
    sc = crt.builtin_scene("cornell", 0)
    W, H = 1024, 1024                      # exactly the maximum: accepted
    assert W * H == crt.PRIMARY_STATIC_MAX_PIXELS
    r = _renderer(crt, sc, W, H)
    _same(r.primary_ids_static(SEED, 0), r.primary_ids(SEED, 0))
    r.close()
    r = _renderer(crt, sc, 1025, 1024)     # one column more: refused with a clear message, crt_primary_ids still serves it
    with pytest.raises(crt.CrtError) as ei:
        r.primary_ids_static(SEED, 0)
    assert ei.value.code == crt.E_LIMIT and "CRT_PRIMARY_STATIC_MAX_PIXELS" in str(ei.value)
    ids, _ = r.primary_ids(SEED, 0)
    assert ids.shape == (1024, 1025)
    r.close()
    sc.free()
