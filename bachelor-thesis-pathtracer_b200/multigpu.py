"""Tile-sharded rendering over the GPUs of one box: one process per GPU, scene and
structures replicated in each GPU's HBM, the frame cut into small tiles dealt round-robin
(`tile mod world`, the generalisation of the reference's 4x4 thread interleave,
src/tracer.cpp:228-231), one NCCL sum-reduce of the radiance buffer to rank 0 per output frame.
Pixels a rank does not own stay exactly zero, so the sum is exact and order independent:
N-GPU image == 1-GPU image bit for bit.

torch.distributed is plumbing only (process group + the collective); the ownership rule itself
lives in libptb (ptb_params.rank/world/tile_w/tile_h) and is mirrored by `owned_mask` for tests.
"""
import numpy as np

TILE_W, TILE_H = 32, 8


def owned_mask(width, height, rank, world, tile_w=TILE_W, tile_h=TILE_H):
    """bool[H, W]: pixels of `rank` — tile (tx, ty) belongs to rank (ty*tiles_x + tx) % world
    (the rule of alloc_frame in csrc/ptb_api.cu)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    tiles_x = (width + tile_w - 1) // tile_w
    ty = np.arange(height)[:, None] // tile_h
    tx = np.arange(width)[None, :] // tile_w
    return ((ty * tiles_x + tx) % world) == rank


class _DevImage:
    """Zero-copy view of libptb's float4[W*H] radiance buffer for torch (CUDA array interface)."""

    def __init__(self, ptr, height, width):
        self.__cuda_array_interface__ = {"shape": (height, width, 4), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def device_image_tensor(pt):
    """torch view (no copy) of the context's radiance buffer in HBM."""
    import torch
    ptr = pt.image_device_ptr()
    if not ptr:
        raise RuntimeError("no frame allocated: call reset()/render() first")
    return torch.as_tensor(_DevImage(ptr, pt.params.height, pt.params.width), device=f"cuda:{pt.device}")


def reduce_image(image, dst=0, group=None):
    """Sum-reduce the per-rank radiance buffers to `dst` (NCCL over NVLink on GPUs, gloo in the CPU
    tests).  In place on `dst`; other ranks' buffers are left unspecified by the collective, so the
    caller must not reuse them — pass a clone when the render continues afterwards."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return image
    dist.reduce(image, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return image


def configure_rank(pt, rank, world, tile_w=TILE_W, tile_h=TILE_H, **params):
    """ptb_set_params with this rank's share of the frame."""
    pt.set_params(rank=rank, world=world, tile_w=tile_w, tile_h=tile_h, **params)


# ---- sample-range sharding (opt-in) -----------------------------------------------------------------
# When 1/N of the frame leaves a GPU fewer pixels than lanes, tile sharding starves it: the samples of a
# pixel are serial in the reference (RNG stream + order-dependent running mean), so pixels bound the
# parallelism.  Sample-range sharding gives every rank ALL pixels and spp/N of the samples from its own
# seeds (ptb_params.seed_offset = rank; rank 0 keeps the reference's seeds).  The result is the same
# estimator with different random numbers: parity is statistical (per-pixel RMSE), not bitwise.
def sample_share(spp, rank, world):
    """Samples rank `rank` takes out of `spp` (differences of at most one)."""
    return spp // world + (1 if rank < spp % world else 0)


def combine_sample_shards(image, n_local, n_total, dst=0, group=None):
    """Each rank holds the reference's running mean over its n_local samples, sum(L)/(n_local+1)
    (tracer.cpp:147-154).  Rebuild sum(L), sum-reduce, and apply the same law over all samples:
    sum(L)/(n_total+1).  In place; the result is valid on `dst`."""
    import torch.distributed as dist
    image.mul_(float(n_local + 1))
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(image, dst=dst, op=dist.ReduceOp.SUM, group=group)
    image.div_(float(n_total + 1))
    image[..., 3] = 1.0
    return image
