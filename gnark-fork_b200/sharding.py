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


def _wait(handle):
    if handle is not None:
        handle.wait()


def compute_h_distributed(engine, exchange, a, b, c, tmp):
    """computeH (prove.go:353-396) over 2^lg ranks.  a, b, c: this rank's contiguous block of the
    zero-padded solver vectors (n / ranks elements each, natural order); tmp: one scratch buffer of
    the same size, or three (one per vector: lets the all-to-all of one vector overlap the phases
    of the next).  exchange(send, recv) is the all-to-all (equal contiguous chunks, chunk q to rank
    q, received chunks ordered by source rank); it returns None when it has completed, or a handle
    whose wait() orders the following phases after it.  Returns the buffer that holds this rank's
    block of h (bit-reversed order, positions [rank n / ranks, (rank+1) n / ranks)): it is `a`."""
    P = api
    vs = (a, b, c)
    ts = tuple(tmp) if isinstance(tmp, (list, tuple)) else (tmp, tmp, tmp)
    shared_tmp = ts[0] is ts[1]
    scales = (P.SCALE_H_A, P.SCALE_H_BC, P.SCALE_H_A)
    pend = [None, None, None]

    def xchg(i, send, recv):
        pend[i] = exchange(send, recv)

    def done(i):
        _wait(pend[i])
        pend[i] = None

    if shared_tmp:                                        # one scratch buffer: vector after vector
        for i, v in enumerate(vs):
            t = ts[0]
            engine.phase(v, t, P.DIST_PACK, 0); xchg(i, t, v); done(i)
            engine.phase(v, t, P.DIST_DIF_LOCAL, 1); xchg(i, v, t); done(i)
            engine.phase(v, t, P.DIST_DIF_CROSS, 1, scales[i])
            engine.phase(v, t, P.DIST_DIT_LOCAL, 0); xchg(i, t, v); done(i)
            engine.phase(v, t, P.DIST_DIT_CROSS, 0)
    else:                                                 # software pipeline over the three vectors
        for i, (v, t) in enumerate(zip(vs, ts)):          # block layout -> cyclic layout
            engine.phase(v, t, P.DIST_PACK, 0)
            xchg(i, t, v)
        for i, (v, t) in enumerate(zip(vs, ts)):          # FFTInverse (DIF) ...
            done(i)
            engine.phase(v, t, P.DIST_DIF_LOCAL, 1)
            xchg(i, v, t)
        for i, (v, t) in enumerate(zip(vs, ts)):          # ... fused with the coset scaling; then FFT (DIT) on the coset
            done(i)
            engine.phase(v, t, P.DIST_DIF_CROSS, 1, scales[i])
            engine.phase(v, t, P.DIST_DIT_LOCAL, 0)
            xchg(i, t, v)
        for i, (v, t) in enumerate(zip(vs, ts)):
            done(i)
            engine.phase(v, t, P.DIST_DIT_CROSS, 0)
    engine.mul_sub(a, b, c)                               # (a*b - c) * den, den carried by H_A
    engine.phase(a, ts[0], P.DIST_DIF_LOCAL, 1)           # FFTInverse (DIF) on the coset
    _wait(exchange(a, ts[0]))
    engine.phase(a, ts[0], P.DIST_DIF_CROSS, 1, P.SCALE_COSETINV_BR)
    return a


def torch_exchange(dist, async_op=False):
    """all_to_all_single over torch.distributed (nccl on CUDA tensors, gloo on CPU tensors).
    On CUDA the collective is ordered after the work already enqueued on torch's current stream and
    wait() orders that stream after the collective: give the library the same stream
    (Context.use_stream(torch.cuda.current_stream().cuda_stream)) and no host synchronisation is
    needed anywhere in computeH."""
    def exchange(send, recv):
        work = dist.all_to_all_single(recv.view(-1), send.view(-1), async_op=async_op)
        return work if async_op else None
    return exchange
