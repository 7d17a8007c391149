"""Host logic of the tile-sharded N>1 path on CPU: world_size-2 gloo processes each take the
pixels libptb would assign them (owned_mask mirrors alloc_frame's rule), fill only those from the
oracle's full-frame image, and the sum-reduce to rank 0 must reproduce the full frame bit for bit
(non-owned pixels are exact zeros)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_owned_masks_partition_the_frame(ptb):
    mg = ptb.multigpu
    for (W, H, world, tw, th) in [(64, 64, 2, 32, 8), (100, 36, 3, 16, 4), (1920, 1080, 8, 32, 8), (8, 8, 4, 32, 8)]:
        total = np.zeros((H, W), np.int32)
        for r in range(world):
            total += mg.owned_mask(W, H, r, world, tw, th)
        assert (total == 1).all()
    # balance at the BASELINE resolution: every rank within 1 % of the mean
    counts = [int(mg.owned_mask(1920, 1080, r, 8).sum()) for r in range(8)]
    assert max(counts) - min(counts) <= 0.01 * np.mean(counts)
    with pytest.raises(ValueError):
        mg.owned_mask(8, 8, 2, 2)


def _worker(rank, world, port, ptbs, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import ptb_b200
    from oracle_lib import MODE_BVH, Oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    W, H = 64, 48
    o = Oracle(ptbs)
    o.build()
    o.configure(W, H, MODE_BVH, nee=1, max_depth=3)
    full, _, _, _ = o.render(2)
    o.close()
    mask = ptb_b200.multigpu.owned_mask(W, H, rank, world, 16, 4)
    mine = np.where(mask[..., None], full, np.float32(0.0)).astype(np.float32)
    t = torch.from_numpy(mine.copy())
    ptb_b200.multigpu.reduce_image(t, dst=0)
    if rank == 0:
        ok = (t.numpy().view(np.uint32) == full.view(np.uint32)).all()
        open(out, "w").write("ok" if ok else "mismatch")
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_reduce_reproduces_full_frame(tmp_path, cornel_ptbs, ptb):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "result.txt")
    mp.spawn(_worker, args=(2, port, cornel_ptbs, out), nprocs=2, join=True)
    assert open(out).read() == "ok"


def _worker_samples(rank, world, port, ptbs, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import ptb_b200
    from oracle_lib import MODE_BVH, Oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mg = ptb_b200.multigpu
    W, H, spp = 48, 32, 7
    o = Oracle(ptbs)
    o.build()
    o.configure(W, H, MODE_BVH, nee=1, max_depth=4)
    yy, xx = np.mgrid[0:H, 0:W]
    seeds = ((xx.astype(np.uint32) << 16) | yy.astype(np.uint32)).reshape(-1)

    def shard(r):  # what rank r renders: all pixels, its share of the samples, seeds offset by r (ptb_params.seed_offset)
        n = mg.sample_share(spp, r, world)
        rng = (seeds + np.uint32((r * 0x9E3779B9) & 0xFFFFFFFF)).astype(np.uint32)
        img, _, _, _ = o.render(n, state=(np.zeros((H, W, 4), np.float32), np.zeros(W * H, np.int32), rng))
        return img.copy(), n

    mine, n_local = shard(rank)
    t = torch.from_numpy(mine.copy())
    mg.combine_sample_shards(t, n_local, spp, dst=0)
    if rank == 0:
        total = np.zeros((H, W, 4), np.float64)
        for r in range(world):
            img, n = shard(r)
            total += img.astype(np.float64) * (n + 1)
        want = total / (spp + 1)
        got = t.numpy().astype(np.float64)
        ok = np.allclose(got[..., :3], want[..., :3], rtol=1e-5, atol=1e-6) and (got[..., 3] == 1.0).all()
        # rank 0 alone reproduces the reference's own first samples (seed_offset 0)
        ref, _, _, _ = o.render(n_local)
        ok = ok and (ref.view(np.uint32) == mine.view(np.uint32)).all()
        open(out, "w").write("ok" if ok else "mismatch")
    o.close()
    dist.barrier()
    dist.destroy_process_group()


def test_sample_range_sharding_combines_to_the_n_plus_1_law(tmp_path, cornel_ptbs, ptb):
    import torch.multiprocessing as mp
    assert [ptb.multigpu.sample_share(7, r, 2) for r in range(2)] == [4, 3]
    assert sum(ptb.multigpu.sample_share(64, r, 8) for r in range(8)) == 64
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "result.txt")
    mp.spawn(_worker_samples, args=(2, port, cornel_ptbs, out), nprocs=2, join=True)
    assert open(out).read() == "ok"
