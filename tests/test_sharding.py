"""N>1 path on CPU: two gloo ranks shard the sample indices, render their share (the oracle stands in
for the device renderer here), merge with the product's one-reduce-per-frame logic, and the merged
sums equal a single-rank render of all samples."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from montecarlopathtracing_b200 import scenes, sharding
from oracle import oracle_api as O

W, H, SPP, DEPTH = 24, 16, 7, 6


def _sums(r):
    n = len(r["counts"])
    s = np.zeros((7, n))
    s[:3] = r["sum"].T
    s[3:6] = r["sumsq"].T
    s[6] = r["counts"]
    return s


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = scenes.load("cbox")
    acc = torch.zeros(7, W * H, dtype=torch.float64)
    for frame_block in range(2):                          # two merges: non-root ranks must contribute deltas only
        first, count = sharding.sample_range(rank, world, SPP)
        r = O.render(sc.camera, W, H, sc.nodes, sc.triangles, sc.materials, DEPTH, 1 << 20, count, seed_mode=1,
                     run_seed=3, first_sample=first + frame_block * SPP)
        acc += torch.from_numpy(_sums(r))
        sharding.merge_sums(acc, 0)
    if rank == 0:
        np.save(out, acc.numpy())
    else:
        assert float(acc.abs().sum()) == 0.0
    dist.barrier()
    dist.destroy_process_group()


def test_sample_range_partitions():
    for world in (1, 2, 3, 8):
        for spp in (0, 1, 7, 8, 4096):
            parts = [sharding.sample_range(r, world, spp) for r in range(world)]
            assert parts[0][0] == 0 and sum(c for _, c in parts) == spp
            for (f0, c0), (f1, _) in zip(parts[:-1], parts[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1


def test_two_rank_merge_equals_single_rank(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "merged.npy")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    merged = np.load(out)
    sc = scenes.load("cbox")
    full = O.render(sc.camera, W, H, sc.nodes, sc.triangles, sc.materials, DEPTH, 1 << 20, 2 * SPP, seed_mode=1, run_seed=3)
    ref = _sums(full)
    assert np.array_equal(merged[6], ref[6])
    np.testing.assert_allclose(merged, ref, rtol=1e-12, atol=0)
    img = sharding.image_from_sums(merged, W, H)
    assert img.shape == (H, W, 4) and np.isfinite(img).all()
    assert (sharding.sigma_of_mean(merged) >= 0).all()
