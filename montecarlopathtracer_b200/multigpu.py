"""spp-split rendering over several GPUs of one box: one process per GPU, disjoint sample
ranges of the same pixels, ONE sum-reduce of the accumulation buffers (NCCL over NVLink) to rank 0.

The reference is single-GPU (CUTracer.cu:222 ``cudaSetDevice(0)``); its only cross-sample coupling is the
running mean of rtx.hlsl:401-404, i.e. a sum, which is what makes the split exact: Philox counters carry the
ABSOLUTE sample index, so the union of the ranks' samples is the 1-GPU sample set.
"""
from __future__ import annotations

import numpy as np


def split_samples(total_spp: int, world: int) -> list[tuple[int, int]]:
    """Strong split of [0, total_spp) into `world` contiguous (offset, count) ranges, sizes differing by <= 1."""
    if total_spp < 0 or world <= 0:
        raise ValueError("total_spp >= 0 and world > 0 required")
    base, rem = divmod(total_spp, world)
    out, off = [], 0
    for r in range(world):
        n = base + (1 if r < rem else 0)
        out.append((off, n))
        off += n
    return out


def rank_samples(rank: int, world: int, spp: int, scaling: str = "weak") -> tuple[int, int]:
    """(sample_offset, count) of one rank. weak: every rank renders `spp` samples (job total = spp*world);
    strong: the job's `spp` samples are divided among the ranks."""
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    if scaling == "weak":
        return rank * spp, spp
    if scaling == "strong":
        return split_samples(spp, world)[rank]
    raise ValueError("scaling must be 'weak' or 'strong'")


def reduce_accumulators(acc, dist=None, dst: int = 0):
    """The path's one exchange step: sum the per-rank (H, W, 4) float32 buffers {sum r, g, b, n} onto `dst`."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(acc, dst=dst, op=dist.ReduceOp.SUM)
    return acc


def finalize(acc) -> np.ndarray:
    """(H, W, 4) sums -> (H, W, 3) mean linear radiance (the `hostcolor` layout of CUTracer.cu:340)."""
    a = acc.detach().cpu().numpy() if hasattr(acc, "detach") else np.asarray(acc)
    n = np.maximum(a[..., 3:4], 1.0)
    return (a[..., :3] / n).astype(np.float32)


def render_distributed(T, cfg_kwargs: dict, width: int, height: int, spp: int, depth: int, rank: int, world: int,
                       dist=None, scaling: str = "strong", device_index: int | None = None):
    """Renders this rank's share on its GPU through libmcpt and reduces; returns the mean image on rank 0
    (None elsewhere).  `T` is montecarlopathtracer_b200.tracer with the scene already loaded."""
    import torch
    off, n = rank_samples(rank, world, spp, scaling)
    acc = torch.zeros(height, width, 4, device="cuda" if device_index is None else f"cuda:{device_index}", dtype=torch.float32)
    if n > 0:
        kw = dict(cfg_kwargs)
        kw["sample_offset"] = kw.get("sample_offset", 0) + off
        cfg = T.render_cfg(width, height, n, depth, **kw)
        T.render_accumulate_device(cfg, acc.data_ptr(), 0, torch.cuda.current_stream().cuda_stream)
    reduce_accumulators(acc, dist)
    return finalize(acc) if rank == 0 else None
