"""Sample-parallel sharding of a frame over ranks.

The reference is single-GPU (no NCCL, no cudaSetDevice anywhere in CudaRayTracer/src). Every (pixel, sample)
of its render kernel is independent (reference: src/CUDAKernels.h:158-162), so a frame shards along the
sample axis: rank g renders spp_g samples of EVERY pixel from the cuRAND subsequences g*W*H + pixel (same
seed, disjoint streams; rank 0 alone reproduces the reference's single-GPU streams), accumulates
float4 {sum_r, sum_g, sum_b, n} per pixel, and ONE reduce(sum) of that buffer to rank 0 finishes the frame;
rank 0 tonemaps. This file holds that host logic only (plain arithmetic + one torch.distributed call); it
runs on CPU tensors with the gloo backend in the tests and on device tensors with NCCL in bench.py.
"""


def shard_spp(spp, world, rank):
    """Samples per pixel rendered by `rank`: spp split as evenly as possible, earlier ranks take the remainder."""
    if world < 1 or not (0 <= rank < world) or spp < 0:
        raise ValueError("bad sharding request spp=%r world=%r rank=%r" % (spp, world, rank))
    return spp // world + (1 if rank < spp % world else 0)


def subsequence_base(rank, width, height):
    """First cuRAND subsequence of `rank`: subsequence = base + pixel (crt_init_rng's second argument)."""
    return int(rank) * int(width) * int(height)


def reduce_accum(accum, dist=None, dst=0):
    """Sum the float4 accumulation buffers of all ranks onto rank `dst` (in place). No-op without a process group."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum


def accum_as_tensor(torch, device_ptr, height, width, device_index):
    """Borrow the library's device accumulation buffer (crt_get_accum's out_dev) as a torch tensor, no copy."""
    class _Holder:
        pass
    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (height, width, 4), "typestr": "<f4", "data": (int(device_ptr), False), "version": 3}
    return torch.as_tensor(h, device=torch.device("cuda", device_index))
