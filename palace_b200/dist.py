"""Multi-GPU plumbing of the `eref` path: one process per GPU, the reference database sharded by
record, and ONE collective — an allgather of fixed-size per-record results.

Records are {ordinal, frag, el, ref_len} as 4 x 32-bit words (bin/extract_ref.cpp:617 prints them).
Shards hold different numbers of records, so each rank pads its block to the largest shard; ordinals
are local to a shard and are offset by the record counts of the shards before it, which restores
the file-order numbering of read_index (:887, :1175).
"""
from __future__ import annotations

import numpy as np

from .eref import RESULT_DTYPE


def pad_records(rec: np.ndarray, slot: int) -> np.ndarray:
    out = np.zeros(slot, RESULT_DTYPE)
    out[: rec.size] = rec
    return out


def merge_gathered(flat: np.ndarray, counts, slot: int) -> np.ndarray:
    """flat: world*slot records as gathered; counts[r]: real records of rank r."""
    parts, base = [], 0
    for r, n in enumerate(counts):
        p = flat[r * slot: r * slot + n].copy()
        p["ordinal"] += base
        base += n
        parts.append(p)
    return np.concatenate(parts) if parts else np.zeros(0, RESULT_DTYPE)


def agree_slot(n_local: int, device) -> int:
    """largest shard's record count: one scalar all_reduce(MAX).  A property of how the database was sharded, so
    a caller that screens many samples against the same shards agrees on it once."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([n_local], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return int(t.item())


def allgather_records(local, n_local: int, slot: int | None = None, compact: bool = True):
    """local: (n_local, 4) int32 tensor on this rank's device (cuda under NCCL, cpu under gloo).
    Returns the merged (sum n, 4) int64 tensor on every rank, ordinals offset into file order.
    The count of each shard travels in the same collective (one extra header row per block) and the merge is
    done on the device without reading the counts back, so the step has ONE collective.
    compact=False: return (body, valid) — the (world, slot, 4) records with global ordinals and the mask of the real
    ones — without squeezing the padding out: boolean indexing must read the number of survivors back, i.e. it makes
    the host wait for the whole step, which a caller that keeps the records on the device does not need."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size()
    if slot is None:
        slot = agree_slot(n_local, local.device)
    if n_local > slot:
        raise ValueError(f"shard holds {n_local} records, agreed slot is {slot}")
    block = torch.zeros((slot + 1, 4), dtype=torch.int32, device=local.device)
    block[0, 0] = n_local
    if n_local:
        block[1: n_local + 1] = local[:n_local]
    out = torch.empty((world * (slot + 1), 4), dtype=torch.int32, device=local.device)
    dist.all_gather_into_tensor(out, block)
    out = out.view(world, slot + 1, 4).to(torch.int64)
    counts = out[:, 0, 0]
    base = torch.cumsum(counts, 0) - counts                      # records of the shards before each one
    body = out[:, 1:, :].clone()
    body[:, :, 0] += base[:, None]
    valid = torch.arange(slot, device=local.device)[None, :] < counts[:, None]
    if not compact:
        return body, valid
    return body[valid]
