"""Multi-GPU plumbing above the C ABI: MSM sharding by point range (one process per GPU,
torch.distributed), partial sums gathered and added on the host -- the GPU analogue of the
reference running its five MultiExp goroutines (prove.go:196-301); no data-path collective.
"""
from __future__ import annotations

import numpy as np

from . import api

FIELDS = (("ar", 12), ("bs1", 12), ("krs", 12), ("krs2", 12), ("bs2", 24))


def shard_range(length: int, index: int, count: int):
    """Same split as groth16.cu:shard_range -- contiguous, balanced, covers [0, length)."""
    return length * index // count, length * (index + 1) // count


def pack(out_dict) -> np.ndarray:
    return np.concatenate([np.asarray(out_dict[k], dtype=np.uint64).reshape(w) for k, w in FIELDS])


def unpack(flat: np.ndarray) -> dict:
    out, o = {}, 0
    for k, w in FIELDS:
        out[k] = flat[o:o + w].copy()
        o += w
    return out


def combine(parts) -> dict:
    """Add per-shard partial sums (G1Jac.AddAssign / G2Jac.AddAssign on the host)."""
    acc = unpack(np.asarray(parts[0], dtype=np.uint64))
    for p in parts[1:]:
        q = unpack(np.asarray(p, dtype=np.uint64))
        for k, w in FIELDS:
            acc[k] = api.g1_jac_add(acc[k], q[k]) if w == 12 else api.g2_jac_add(acc[k], q[k])
    return acc


def gather_combine(dist, out_dict, device=None):
    """all_gather the 72-word partial results of every rank; every rank returns the combined
    Jacobian points.  `dist` is torch.distributed (nccl with device=cuda:k, or gloo on CPU)."""
    import torch
    flat = torch.from_numpy(pack(out_dict).view(np.int64).copy())
    if device is not None:
        flat = flat.to(device)
    gathered = [torch.empty_like(flat) for _ in range(dist.get_world_size())]
    dist.all_gather(gathered, flat)
    return combine([g.cpu().numpy().view(np.uint64) for g in gathered])
