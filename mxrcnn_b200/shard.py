"""Image sharding across the GPUs of one box.

Every function of the hot path is per-image (the reference enforces batch 1 and puts one image on each
device, rcnn/loader.py:273-281), so images shard with no halo and no collective in the data path:
image i of a global batch goes to rank ``i % world``.  The only exchange is the harness's final
all-gather of the fixed-shape per-image outputs (NCCL over NVLink on GPUs, gloo in the CPU tests).
torch.distributed is imported lazily: the library itself never needs it.
"""
import numpy as np


def shard_indices(n_images, rank, world):
    """Global image indices owned by ``rank`` (round-robin, ascending)."""
    return list(range(rank, n_images, world))


def unshard(per_rank_rows, n_images, world):
    """Inverse of the round-robin split: ``per_rank_rows[r][j]`` is the result of global image
    ``r + j*world``.  Returns the list of per-image results in global image order."""
    out = [None] * n_images
    for r in range(world):
        for j, i in enumerate(shard_indices(n_images, r, world)):
            out[i] = per_rank_rows[r][j]
    return out


def all_gather_outputs(local, n_images, group=None):
    """All-gather per-image fixed-shape outputs.

    ``local``: tensor (n_local, ...) holding this rank's images in ascending global order, padded to
    ``ceil(n_images / world)`` rows if the split is uneven.  Returns a tensor (n_images, ...) in global
    image order, identical on every rank and independent of the world size.
    """
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local[:n_images]
    per = (n_images + world - 1) // world
    if local.shape[0] < per:
        pad = torch.zeros((per - local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        local = torch.cat([local, pad])
    parts = [torch.empty_like(local) for _ in range(world)]
    dist.all_gather(parts, local.contiguous(), group=group)
    rows = unshard([p for p in parts], n_images, world)
    return torch.stack(rows)


def checksum(keep_rows):
    """Order-sensitive checksum of keep lists (a checksum of checksums: one per image, then combined)."""
    M = (1 << 64) - 1
    acc = 1469598103934665603
    for row in keep_rows:
        h = 1469598103934665603
        for v in np.asarray(row, dtype=np.int64).ravel().tolist():
            h = ((h ^ (v & 0xFFFFFFFF)) * 1099511628211) & M
        acc = ((acc ^ h) * 1099511628211) & M
    return acc
