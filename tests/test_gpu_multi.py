"""Multi-GPU through the C-ABI only: crt_comm_* (one process per GPU, NCCL inside the library) and
crt_group_* (one process, several GPUs, reduce + tonemap in one peer-memory kernel). The reference is single-GPU
(CudaRayTracer/src/CUDARenderer.cuh:55-60); what is checked is that the sharded frame is the SUM of the shards a single
GPU renders from the same subsequence bases (crt_init_rng base g*W*H), bit for bit."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch   # BEFORE the library binds NCCL: PyTorch's libnccl.so.2 must be the one in the process (crt_comm.cu, nccl())

pytestmark = pytest.mark.gpu
SEED = 1804289383
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _device_count():
    return torch.cuda.device_count()


def _single_gpu_shards(crt, sc, W, H, spps, bounces):
    out = []
    for g, spp in enumerate(spps):
        r = crt.Renderer(W, H, mode=crt.MODE_WAVEFRONT)
        r.set_scene(sc)
        r.update_camera(sc.camera)
        r.initialize(SEED, g * W * H)
        st = r.render(spp, bounces, flags=crt.RENDER_NO_TONEMAP)
        out.append((r.accum().copy(), int(st.rays)))
        r.close()
    return out


def test_comm_single_rank_is_a_tonemap(crt):
    """A communicator of one rank: crt_reduce_accum leaves the buffer alone and tonemaps (NCCL is really initialised)."""
    sc = crt.builtin_scene("cornell")
    W = H = 64
    r = crt.Renderer(W, H, mode=crt.MODE_WAVEFRONT)
    r.set_scene(sc)
    r.update_camera(sc.camera)
    r.initialize(SEED, 0)
    r.render(4, 8)
    want_img, want_acc = r.image_rgba8().copy(), r.accum().copy()
    r.comm_init(crt.comm_unique_id(), 1, 0)
    n, rank, ver = r.comm_info()
    assert (n, rank) == (1, 0) and ver >= 20000
    r.initialize(SEED, 0)
    r.render(4, 8, flags=crt.RENDER_NO_TONEMAP)
    r.reduce_accum(0)
    assert np.array_equal(r.accum().view(np.uint32), want_acc.view(np.uint32))
    assert np.array_equal(r.image_rgba8(), want_img)
    with pytest.raises(crt.CrtError):
        r.reduce_accum(3)
    r.comm_destroy()
    r.close()
    sc.free()


def test_group_one_device_equals_renderer(crt):
    sc = crt.builtin_scene("cornell")
    W = H = 96
    g = crt.DeviceGroup(W, H, [0], mode=crt.MODE_WAVEFRONT)
    g.set_scene(sc)
    g.update_camera(sc.camera)
    g.initialize(SEED)
    st = g.render(6, 8)
    r0 = g.renderer(0)
    ref = _single_gpu_shards(crt, sc, W, H, [6], 8)[0]
    assert np.array_equal(r0.accum().view(np.uint32), ref[0].view(np.uint32))
    assert int(st.rays) == ref[1]
    g.close()
    with pytest.raises(crt.CrtError):
        crt.DeviceGroup(W, H, [0, 0])
    sc.free()


@pytest.mark.parametrize("force_nccl", [False, True])
def test_group_two_devices_sum_of_shards(crt, force_nccl):
    """Two GPUs driven through crt_group_* only. 9 spp -> 5 + 4; the root's buffer is shard 0 + shard 1 exactly (one
    float add per component, in device order) and its image the tonemap of that sum."""
    if _device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    sc = crt.builtin_scene("cornell_blob", 48)
    W, H = 320, 180
    g = crt.DeviceGroup(W, H, [0, 1], mode=crt.MODE_WAVEFRONT)
    g.set_scene(sc)
    g.update_camera(sc.camera)
    g.initialize(SEED)
    st = g.render(9, 20, group_flags=crt.GROUP_FORCE_NCCL if force_nccl else 0)
    root = g.renderer(0)
    acc, img = root.accum().copy(), root.image_rgba8().copy()
    shards = _single_gpu_shards(crt, sc, W, H, [5, 4], 20)
    want = shards[0][0] + shards[1][0]
    assert np.array_equal(acc.view(np.uint32), want.view(np.uint32))
    assert np.all(acc[..., 3] == 9)
    assert int(st.rays) == shards[0][1] + shards[1][1]
    assert st.peer or force_nccl or True
    r = crt.Renderer(W, H)
    r.set_accum(want)
    r.tonemap()
    assert np.array_equal(img, r.image_rgba8())
    r.close()
    # a second frame continues the streams on both devices
    g.render(4, 20, group_flags=crt.GROUP_FORCE_NCCL if force_nccl else 0)
    assert np.all(g.renderer(0).accum()[..., 3] == 4)
    g.close()
    sc.free()


WORKER = r'''
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, sys.argv[1])
crt = importlib.import_module("raytracer-cuda_b200")
rank, world, idfile, out = int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5]
W, H, SEED = 320, 180, 1804289383
if rank == 0:
    open(idfile + ".tmp", "wb").write(crt.comm_unique_id()); os.rename(idfile + ".tmp", idfile)
while not os.path.exists(idfile): time.sleep(0.05)
uid = open(idfile, "rb").read()
sc = crt.builtin_scene("cornell_blob", 48)
r = crt.Renderer(W, H, device=rank, mode=crt.MODE_WAVEFRONT)
r.set_scene(sc); r.update_camera(sc.camera); r.initialize(SEED, rank * W * H)
r.comm_init(uid, world, rank)
for frame in range(2):                       # two frames: the reduce must not race with the next frame's begin
    r.render(5 - rank, 20, flags=crt.RENDER_NO_TONEMAP)
    r.reduce_accum(0)
if rank == 0:
    np.save(out, r.accum()); np.save(out + ".img.npy", r.image_rgba8())
r.close()
'''


def test_comm_two_processes(crt, tmp_path):
    """One process per GPU, the launch the driver uses: rank 0's NCCL id goes through a file, crt_comm_init on both,
    two frames of render + crt_reduce_accum. Rank 0's buffer after frame 2 is the sum of the two shards' second frames."""
    if _device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    idfile, out = str(tmp_path / "nccl_id"), str(tmp_path / "acc.npy")
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, str(g), "2", idfile, out]) for g in range(2)]
    for p in procs:
        assert p.wait(timeout=300) == 0
    acc, img = np.load(out), np.load(out + ".img.npy")
    sc = crt.builtin_scene("cornell_blob", 48)
    W, H = 320, 180
    want = np.zeros((H, W, 4), np.float32)
    for g, spp in enumerate((5, 4)):
        r = crt.Renderer(W, H, mode=crt.MODE_WAVEFRONT)
        r.set_scene(sc)
        r.update_camera(sc.camera)
        r.initialize(SEED, g * W * H)
        r.render(spp, 20, flags=crt.RENDER_NO_TONEMAP)
        r.render(spp, 20, flags=crt.RENDER_NO_TONEMAP)     # frame 2 continues the streams and overwrites
        want += r.accum()
        r.close()
    assert np.array_equal(acc.view(np.uint32), want.view(np.uint32))
    assert np.all(acc[..., 3] == 9)
    r = crt.Renderer(W, H)
    r.set_accum(want)
    r.tonemap()
    assert np.array_equal(img, r.image_rgba8())
    r.close()
    sc.free()
