"""Multi-GPU plumbing: sample-index sharding and the one-reduce-per-frame merge.

The path shards by sample index: every rank holds a replicated scene + BVH and its own accumulators,
renders a disjoint range of global sample indices (seed_mode 1: the LCG is re-seeded from
(pixel, global sample index), so the union over ranks is the same sample set a single GPU would
draw), and the 7 planes of per-pixel sums (sum rgb, sum of squares rgb, counted samples) are merged
by ONE reduce onto rank 0.  No other collective exists on the path.
"""
import numpy as np


def sample_range(rank, world, spp_total):
    """Contiguous block of global sample indices [first, first + count) owned by `rank`."""
    base, rem = divmod(int(spp_total), int(world))
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


def merge_sums(sums, dst=0, group=None):
    """In-place reduce of the accumulators onto `dst` (torch tensor on the device the backend serves).
    Ranks other than dst clear their buffer afterwards, so the next merge contributes only new samples."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return sums
    dist.reduce(sums, dst=dst, group=group)
    if dist.get_rank(group) != dst:
        sums.zero_()
    return sums


def image_from_sums(sums, width, height):
    """history.cl semantics from mergeable sums: mean of the counted (non-zero) samples per pixel.
    sums: (7, W*H) array-like -> float32 (H, W, 4) with w = 0, row 0 = image bottom."""
    s = np.asarray(sums, dtype=np.float64).reshape(7, -1)
    n = np.maximum(s[6], 1.0)
    img = np.zeros((height, width, 4), np.float32)
    img[..., :3] = (s[:3] / n).T.reshape(height, width, 3)
    return img


def sigma_of_mean(sums):
    """Per-pixel standard error of the mean from the sums (for the 4-sigma parity test)."""
    s = np.asarray(sums, dtype=np.float64).reshape(7, -1)
    n = np.maximum(s[6], 1.0)
    mean = s[:3] / n
    var = np.maximum(s[3:6] / n - mean * mean, 0.0) * (n / np.maximum(n - 1.0, 1.0))
    return np.sqrt(var / n)


def render_sharded(ctx, spp_total, rank=0, world=1, sums_tensor=None, group=None):
    """Render this rank's share of `spp_total` samples with ctx (seed_mode 1) and merge onto rank 0.
    sums_tensor: the device tensor bound with ctx.bind_sums (required when world > 1)."""
    first, count = sample_range(rank, world, spp_total)
    ctx.set_first_sample(first)
    ctx.render(count)
    if world > 1:
        merge_sums(sums_tensor, 0, group)
    return first, count
