"""One-process-per-GPU layout of the checksum pass: who hashes what, and how digests come home.

The path shards with no exchange step: units are independent, so each rank
hashes a byte-balanced share and only the 16-byte digests travel -- over the host, through a gloo
group, never NCCL.  rclone's counterpart is its pool of 8 checker goroutines pulling pairs off a
channel ([rclone] fs/sync pairChecker, reached from
/root/reference/task/common/machine/storage.go:158); here the pool is the set of ranks.

`hasher` is whatever provides ``hash_buffers(base, offsets, lengths) -> uint8[n,16]``: in production
a ``Context`` bound to this rank's GPU.  Nothing here falls back to a CPU hash.
"""
from __future__ import annotations

from typing import Optional

import numpy as np

from .binding import shard_lpt


def my_units(lengths, rank: int, world: int) -> np.ndarray:
    """Indices (ascending) of the units this rank hashes."""
    assign = shard_lpt(lengths, world)
    return np.nonzero(assign == rank)[0].astype(np.int64)


def gather_manifest(local_digests: np.ndarray, local_units: np.ndarray, n_total: int, group=None) -> np.ndarray:
    """All ranks contribute (unit index, digest) rows; every rank gets the full manifest uint8[n_total,16].

    Uses CPU tensors on a gloo group (pass one when the default group is NCCL)."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    counts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([local_units.size], dtype=torch.int64), group=group)
    cap = max(1, max(int(c.item()) for c in counts))
    # one padded row block per rank: 8 bytes of unit index + 16 bytes of digest
    mine = np.zeros((cap, 24), dtype=np.uint8)
    if local_units.size:
        mine[: local_units.size, :8] = local_units.astype("<i8").view(np.uint8).reshape(-1, 8)
        mine[: local_units.size, 8:] = local_digests
    blocks = [torch.zeros((cap, 24), dtype=torch.uint8) for _ in range(world)]
    dist.all_gather(blocks, torch.from_numpy(mine), group=group)
    manifest = np.zeros((n_total, 16), dtype=np.uint8)
    seen = np.zeros(n_total, dtype=bool)
    for c, blk in zip(counts, blocks):
        k = int(c.item())
        if not k:
            continue
        b = blk.numpy()[:k]
        idx = b[:, :8].copy().view("<i8").reshape(-1)
        if seen[idx].any():
            raise RuntimeError("gather_manifest: a unit was hashed by two ranks")
        seen[idx] = True
        manifest[idx] = b[:, 8:]
    if not seen.all():
        raise RuntimeError(f"gather_manifest: {int((~seen).sum())} units were hashed by no rank")
    return manifest


def hash_sharded(hasher, base, offsets, lengths, rank: int, world: int, group=None) -> np.ndarray:
    """Hash this rank's share with `hasher`, then assemble the whole manifest on every rank."""
    off = np.ascontiguousarray(offsets, dtype=np.uint64)
    ln = np.ascontiguousarray(lengths, dtype=np.uint64)
    units = my_units(ln, rank, world)
    local = hasher.hash_buffers(base, off[units], ln[units]) if units.size else np.zeros((0, 16), np.uint8)
    if world == 1:
        return local
    return gather_manifest(local, units, off.size, group=group)
