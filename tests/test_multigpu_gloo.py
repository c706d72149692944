"""N>1 host logic on CPU: world_size-2 gloo process group (no GPU).

Each rank plays one GPU of bench.py's multi-GPU path: it takes its share of the spp budget and its cuRAND
subsequence base from raytracer-cuda_b200/multigpu.py, renders its samples with the CPU ORACLE standing in for
the device (the oracle is test infrastructure, allowed here), and the float4 accumulation buffers are summed
onto rank 0 with the same reduce_accum() call bench.py makes over NCCL. Rank 0's result must equal the sum of
the per-rank oracle renders computed in one process, and rank 0's shard must be the single-GPU stream.
"""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SEED = 1804289383
W, H, SPP, BOUNCES = 48, 32, 5, 8


def _accum_for_rank(crt, ol, mg, rank, world):
    sc = crt.builtin_scene("cornell")
    osc = ol.OracleScene(sc.pointer())
    spp = mg.shard_spp(SPP, world, rank)
    rad, _, rays = osc.render(sc.camera, W, H, spp, BOUNCES, SEED, subseq_base=mg.subsequence_base(rank, W, H))
    acc = np.zeros((H, W, 4), np.float32)
    acc[..., :3] = rad * np.float32(spp)      # the oracle returns the mean; the device buffer holds sums
    acc[..., 3] = spp
    osc.close()
    sc.free()
    return acc, rays


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    crt = importlib.import_module("raytracer-cuda_b200")
    mg = importlib.import_module("raytracer-cuda_b200.multigpu")
    import oracle_lib as ol
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    acc, rays = _accum_for_rank(crt, ol, mg, rank, world)
    t = torch.from_numpy(acc.copy())
    mg.reduce_accum(t, dist, dst=0)
    tot = torch.tensor([float(rays)], dtype=torch.float64)
    dist.all_reduce(tot)
    if rank == 0:
        np.save(os.path.join(out_dir, "reduced.npy"), t.numpy())
        np.save(os.path.join(out_dir, "rays.npy"), tot.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_shard_arithmetic():
    mg = importlib.import_module("raytracer-cuda_b200.multigpu")
    for spp in (0, 1, 7, 2000):
        for world in (1, 2, 3, 8):
            parts = [mg.shard_spp(spp, world, r) for r in range(world)]
            assert sum(parts) == spp and max(parts) - min(parts) <= 1 and parts == sorted(parts, reverse=True)
    assert mg.shard_spp(2000, 8, 3) == 250
    assert mg.subsequence_base(0, 2560, 1440) == 0 and mg.subsequence_base(3, 2560, 1440) == 3 * 3686400
    with pytest.raises(ValueError):
        mg.shard_spp(4, 2, 2)


def test_two_rank_gloo_reduce_matches_single_process_sum(crt, tmp_path):
    import torch.multiprocessing as mp
    mg = importlib.import_module("raytracer-cuda_b200.multigpu")
    import oracle_lib as ol
    world = 2
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    reduced = np.load(tmp_path / "reduced.npy")
    parts = [_accum_for_rank(crt, ol, mg, r, world) for r in range(world)]
    expect = parts[0][0] + parts[1][0]
    assert np.array_equal(reduced.view(np.uint32), expect.view(np.uint32))
    assert np.all(reduced[..., 3] == SPP)
    assert float(np.load(tmp_path / "rays.npy")[0]) == float(parts[0][1] + parts[1][1])
    # rank 0's shard is the reference's single-GPU stream (subsequence = pixel), truncated to its spp
    sc = crt.builtin_scene("cornell")
    osc = ol.OracleScene(sc.pointer())
    single, _, _ = osc.render(sc.camera, W, H, mg.shard_spp(SPP, world, 0), BOUNCES, SEED)
    assert np.array_equal((parts[0][0][..., :3] / parts[0][0][..., 3:4]).view(np.uint32), single.view(np.uint32)) or \
        np.allclose(parts[0][0][..., :3] / parts[0][0][..., 3:4], single, rtol=1e-6)
    # the two ranks drew from disjoint streams: their estimates differ
    assert not np.array_equal(parts[0][0][..., :3] / 3.0, parts[1][0][..., :3] / 2.0)
    osc.close()
    sc.free()
