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


# ------------------------------------------------------------------------------------------
# Multi-GPU PlonK (P1 range-sharded commitments, P4 coset-parallel quotient) over the phase
# ABI b200_plonk_session_*: every rank holds its shard of the key (plonk_pk_upload(shard_index, shard_count)),
# runs the same phases on the same solution and blinding, and between phases
#   * the partial commitments (96-byte Jacobian points) are all-gathered and added on the host, and
#   * after phase 2 the numerator's coset blocks are all-gathered in place on the device (coset k, computed by
#     rank k mod R, is block bitrev(k) of the bit-reversed big-domain buffer: prove.go:1029-1035).
# ------------------------------------------------------------------------------------------
def coset_block(k: int, log_blocks: int) -> int:
    """Block of the numerator buffer that holds coset k: bitrev_{log_big}(rho i + k) = bitrev(k) n + bitrev_{log_n}(i)."""
    r = 0
    for b in range(log_blocks):
        r |= ((k >> b) & 1) << (log_blocks - 1 - b)
    return r


def sum_partials(parts) -> np.ndarray:
    """parts: per-rank (m, 12) partial G1Jac commitments -> (m, 12) totals."""
    acc = np.array(parts[0], dtype=np.uint64).reshape(-1, 12).copy()
    for p in parts[1:]:
        q = np.asarray(p, dtype=np.uint64).reshape(-1, 12)
        for j in range(acc.shape[0]):
            acc[j] = api.g1_jac_add(acc[j], q[j])
    return acc


def _gather_points(dist, partials: np.ndarray, device=None) -> np.ndarray:
    import torch
    flat = torch.from_numpy(np.ascontiguousarray(partials, dtype=np.uint64).view(np.int64).reshape(-1).copy())
    if device is not None:
        flat = flat.to(device)
    gathered = [torch.empty_like(flat) for _ in range(dist.get_world_size())]
    dist.all_gather(gathered, flat)
    return sum_partials([g.cpu().numpy().view(np.uint64) for g in gathered])


def plonk_exchange_numerator(dist, session, device):
    """All-gather the coset blocks of the quotient numerator in place (NCCL over NVLink)."""
    import torch
    ptr, block_elems, log_blocks = session.numerator()
    rho, world, rank = 1 << log_blocks, dist.get_world_size(), dist.get_rank()
    words = block_elems * 4

    class _Mem:      # __cuda_array_interface__ view of the library's buffer, no copy
        def __init__(self, p, n):
            self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i8", "data": (p, False), "version": 2}
    buf = torch.as_tensor(_Mem(ptr, rho * words), device=device)
    mine = [k for k in range(rho) if k % world == rank]
    rounds = (rho + world - 1) // world
    for j in range(rounds):                      # round j moves coset j * world + r of every rank r that has one
        owners = [r for r in range(world) if j * world + r < rho]
        if len(owners) == world:
            outs = [buf[coset_block(j * world + r, log_blocks) * words:][:words] for r in range(world)]
            dist.all_gather(outs, outs[rank])
        else:                                     # fewer cosets than ranks in the last round: broadcasts
            for r in owners:
                blk = buf[coset_block(j * world + r, log_blocks) * words:][:words]
                dist.broadcast(blk, src=r)
    torch.cuda.synchronize()                      # the library's stream continues from here
    return mine


def plonk_prove_distributed(dist, ctx, pk, l, r, o, public_witness, blinding, bsb=None, device=None):
    """One PlonK proof over dist.get_world_size() GPUs; every rank returns the same (PlonkProof, timings)."""
    s = ctx.plonk_session(pk, l, r, o, public_witness, blinding, bsb)
    try:
        s.step(1, _gather_points(dist, s.partials, device))
        s.step(2, _gather_points(dist, s.partials, device))
        plonk_exchange_numerator(dist, s, device)
        s.step(3)
        s.step(4, _gather_points(dist, s.partials, device))
        s.step(5, _gather_points(dist, s.partials, device))
        return s.finish(_gather_points(dist, s.partials, device))
    except Exception:
        s.abandon()
        raise
