"""DM-trial sharding across ranks and the single gather of candidate lists.

The reference fans DM trials out to one pthread per GPU of ONE process and merges the workers'
candidate vectors after pthread_join (src/pipeline_multi.cu:342-359).  Here ranks are processes
(one per GPU, torch.distributed): DM trials are independent units, so the data path has no
collective; the only exchange is one gather of the flattened candidate records to rank 0, which
restores the canonical ascending-dm_idx order of the reference's `-t 1` run.
Works with the NCCL backend (device tensors over NVLink) and with gloo (CPU tests)."""
from __future__ import annotations

import numpy as np

from .host import CAND_DTYPE


def shard_range(ndm: int, rank: int, world: int):
    """Contiguous block partition of [0, ndm): returns (start, count); sizes differ by <= 1."""
    base, rem = divmod(ndm, world)
    start = rank * base + min(rank, rem)
    return start, base + (1 if rank < rem else 0)


def canonical_order(recs: np.ndarray) -> np.ndarray:
    """Stable sort of flattened records by the dm_idx of their top-level entry: associates stay
    behind their parent, emission order inside a DM trial is preserved."""
    if len(recs) == 0:
        return recs
    top = np.maximum.accumulate(np.where(recs["is_top"] == 1, np.arange(len(recs)), 0))
    key = recs["dm_idx"][top]
    return recs[np.argsort(key, kind="stable")]


def gather_candidates(recs: np.ndarray, device=None, group=None) -> np.ndarray | None:
    """Gather every rank's records on rank 0 (others return None).  One all_gather of the counts
    plus one padded all_gather of the 28-byte records viewed as int32 words."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return canonical_order(recs)
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = device if device is not None else torch.device("cpu")
    words = CAND_DTYPE.itemsize // 4
    n = torch.tensor([len(recs)], dtype=torch.int64, device=dev)
    counts = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    cap = max(max(counts), 1)
    buf = torch.zeros(cap * words, dtype=torch.int32, device=dev)
    if len(recs):
        flat = torch.from_numpy(np.ascontiguousarray(recs).view(np.int32).reshape(-1).copy())
        buf[:len(recs) * words] = flat.to(dev)
    out = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    if rank != 0:
        return None
    parts = [o[:c * words].cpu().numpy().view(CAND_DTYPE) for o, c in zip(out, counts)]
    return canonical_order(np.concatenate(parts))
