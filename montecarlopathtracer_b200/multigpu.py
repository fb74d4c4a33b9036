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


def comm_init_from_torch(T, dist, rank: int, world: int) -> None:
    """Bootstraps libmcpt's own NCCL communicator (mcpt_comm_init) with torch.distributed as the side channel for the
    128-byte id -- the only thing torch.distributed is used for on the multi-GPU path; the reduce itself runs inside
    libmcpt.so (mcpt_render_scene_distributed).  A C++ caller hands the id over by file, socket or MPI instead."""
    if world == 1:
        T.comm_init(0, 1)
        return
    box = [T.comm_get_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    T.comm_init(rank, world, box[0])


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


def render_distributed(T, cfg_kwargs: dict, width: int, height: int, spp: int, depth: int):
    """All ranks call this after comm_init_from_torch: the WHOLE job's cfg goes to mcpt_render_scene_distributed, which
    renders this rank's share, reduces with NCCL inside libmcpt.so and returns the mean image on rank 0 (None elsewhere)."""
    return T.render_scene_distributed(T.render_cfg(width, height, spp, depth, **cfg_kwargs))
