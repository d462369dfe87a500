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


# ------------------------------------------------------------------------------------------
# Distributed computeH: each transform = [local phase] -> all-to-all -> [cross phase]
# (include/b200prover.h: b200_ntt_dist_phase; csrc/ntt.cuh for the layouts).
# ------------------------------------------------------------------------------------------
def z_shard_range(n_z: int, log_n: int, index: int, count: int):
    """Same split as groth16.cu:shard_range_z: with a power-of-two shard count G1.Z follows the block
    layout the distributed computeH leaves h in."""
    n = 1 << log_n
    if count & (count - 1) == 0 and count <= n:
        block = n // count
        return min(n_z, block * index), min(n_z, block * (index + 1))
    return shard_range(n_z, index, count)


def log2_ranks(world: int) -> int:
    lg = world.bit_length() - 1
    if world < 2 or (1 << lg) != world:
        raise ValueError("the distributed computeH needs a power-of-two number of ranks >= 2")
    return lg


class DeviceEngine:
    """The device side of a phase: b200_ntt_dist_phase / b200_fr_mul_sub_dev on torch CUDA tensors
    (the tensors only lend their memory; the kernels are the library's)."""

    def __init__(self, ctx, log_n, generator, coset_gen, rank, world):
        self.ctx, self.log_n, self.gen, self.coset = ctx, log_n, generator, coset_gen
        self.rank, self.lg = rank, log2_ranks(world)

    def phase(self, data, tmp, kind, inverse, scale=api.SCALE_NONE):
        self.ctx.ntt_dist_phase(self.log_n, self.gen, self.coset, data, tmp, self.lg, self.rank, kind, inverse, scale)

    def mul_sub(self, a, b, c):
        self.ctx.fr_mul_sub_dev(a, b, c, (1 << self.log_n) >> self.lg)


def compute_h_distributed(engine, exchange, a, b, c, tmp):
    """computeH (prove.go:353-396) over 2^lg ranks.  a, b, c: this rank's contiguous block of the
    zero-padded solver vectors (n / ranks elements each, natural order); tmp: one scratch buffer of
    the same size.  exchange(send, recv) is the all-to-all (equal contiguous chunks, chunk q to rank
    q, received chunks ordered by source rank).  Returns the buffer that holds this rank's block of h
    (bit-reversed order, positions [rank n / ranks, (rank+1) n / ranks)): it is `a`."""
    P = api
    for v in (a, b, c):                                   # block layout -> cyclic layout
        engine.phase(v, tmp, P.DIST_PACK, 0)
        exchange(tmp, v)
    for v, scale in ((a, P.SCALE_H_A), (b, P.SCALE_H_BC), (c, P.SCALE_H_A)):   # FFTInverse (DIF) fused with the coset scaling
        engine.phase(v, tmp, P.DIST_DIF_LOCAL, 1)
        exchange(v, tmp)
        engine.phase(v, tmp, P.DIST_DIF_CROSS, 1, scale)
    for v in (a, b, c):                                   # FFT (DIT) on the coset
        engine.phase(v, tmp, P.DIST_DIT_LOCAL, 0)
        exchange(tmp, v)
        engine.phase(v, tmp, P.DIST_DIT_CROSS, 0)
    engine.mul_sub(a, b, c)                               # (a*b - c) * den, den carried by H_A
    engine.phase(a, tmp, P.DIST_DIF_LOCAL, 1)             # FFTInverse (DIF) on the coset
    exchange(a, tmp)
    engine.phase(a, tmp, P.DIST_DIF_CROSS, 1, P.SCALE_COSETINV_BR)
    return a


def torch_exchange(dist):
    """all_to_all_single over torch.distributed (nccl on CUDA tensors, gloo on CPU tensors)."""
    def exchange(send, recv):
        dist.all_to_all_single(recv.view(-1), send.view(-1))
        if recv.is_cuda:
            import torch
            torch.cuda.current_stream(recv.device).synchronize()   # the library runs on its own stream
    return exchange
