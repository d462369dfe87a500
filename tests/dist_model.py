"""CPU stand-in for the device side of the distributed computeH (TEST INFRASTRUCTURE): the phases
of gnark-fork_b200/csrc/ntt.cu replayed by tests/cpu_model (same index-map headers), so that the
orchestration and the exchange in gnark-fork_b200/sharding.py run without a GPU."""
import ctypes

import numpy as np

from oracle import bn254 as B
from oracle import ntt as N


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _arr(x):
    """numpy view of a numpy array or a CPU torch tensor (int64 view of u64 limbs)."""
    if hasattr(x, "numpy"):
        return x.numpy().view(np.uint64)
    return x


def bitrev(i, bits):
    return int(format(i, f"0{bits}b")[::-1], 2) if bits else 0


class ModelEngine:
    def __init__(self, model, cref, log_n, rank, world, tl=4, smax=3):
        self.model, self.log_n, self.rank = model, log_n, rank
        self.lg = world.bit_length() - 1
        self.tl, self.smax = tl, smax
        d = N.Domain(1 << log_n)
        n, r = 1 << log_n, B.R
        g, ninv = d.coset_gen, d.cardinality_inv
        den = pow((pow(g, n, r) - 1) % r, r - 2, r)
        ginv = pow(g, r - 2, r)
        self.gen = cref.fr(d.generator)
        br = [bitrev(i, log_n) for i in range(n)]
        self.tables = {
            0: None,
            1: cref.ints_to_mont([pow(g, br[i], r) * ninv % r * den % r for i in range(n)], r),   # H_A
            2: cref.ints_to_mont([pow(g, br[i], r) * ninv % r for i in range(n)], r),             # H_BC
            3: cref.ints_to_mont([pow(ginv, br[i], r) * ninv % r for i in range(n)], r),          # COSETINV_BR
        }

    def phase(self, data, tmp, kind, inverse, scale=0):
        t = self.tables[scale]
        d, s = _arr(data), _arr(tmp)
        self.model.model_ntt_dist_phase(_p(d), _p(s), self.log_n, self.lg, self.rank, int(kind), int(inverse),
                                        self.tl, self.smax, _p(self.gen), _p(t) if t is not None else None)

    def mul_sub(self, a, b, c):
        a_, b_, c_ = _arr(a), _arr(b), _arr(c)
        self.model.model_fr_mul_sub(_p(a_), _p(b_), _p(c_), ctypes.c_size_t(a_.shape[0]))
