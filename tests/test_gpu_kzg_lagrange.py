"""kzg.ToLagrangeG1 on the device (b200_kzg_to_lagrange_g1; plonk Setup, /root/reference/backend/plonk/bn254/setup.go:134)
against (a) the oracle's restatement of the group FFT (oracle/plonk_prove.to_lagrange_g1: points in, points out, no tau)
and (b) the definition: out[i] = [L_i(tau)]G1 for the tau the test SRS was made from -- two independent routes."""
import random

import numpy as np
import pytest

from oracle import bn254 as B
from oracle import ntt as N
from oracle import plonk_prove as PP
from tests import util

pytestmark = pytest.mark.gpu


def _srs(ctx, cref, n, tau, mod=B):
    mono = [pow(tau, j, mod.R) for j in range(n)]
    gen = cref.g1_points_to_array([mod.G1_GEN])[0]
    return mono, ctx.batch_scalar_mul(gen, cref.ints_to_mont(mono, mod.R), 1)


@pytest.mark.parametrize("log_n", [0, 1, 3, 5])
def test_matches_oracle_group_fft(ctx, cref, log_n):
    n = 1 << log_n
    tau = random.Random(100 + log_n).randrange(2, B.R)
    mono, srs = _srs(ctx, cref, n, tau)
    d = N.Domain(n)
    got = ctx.kzg_to_lagrange_g1(srs, cref.fr(d.generator))
    pts = [B.g1_mul(B.G1_GEN, m) for m in mono]
    want = PP.to_lagrange_g1(pts, d)
    assert [tuple(p) for p in map(util.g1_aff_ints, got)] == [tuple(p) for p in want]


@pytest.mark.parametrize("log_n", [4, 10, 13])
def test_matches_lagrange_basis_at_tau(ctx, cref, log_n):
    """[L_i(tau)]G1 from the scalar side: L(tau) = F^-1 (tau^j) (oracle/ntt.py), then the fixed-base kernel."""
    n = 1 << log_n
    tau = random.Random(200 + log_n).randrange(2, B.R)
    mono, srs = _srs(ctx, cref, n, tau)
    d = N.Domain(n)
    lag = list(mono)
    N.fft_inverse(lag, d, N.DIF)
    N.bit_reverse(lag)
    gen = cref.g1_points_to_array([B.G1_GEN])[0]
    want = ctx.batch_scalar_mul(gen, cref.ints_to_mont(lag, B.R), 1)
    got = ctx.kzg_to_lagrange_g1(srs, cref.fr(d.generator))
    assert np.array_equal(got, want)
    # and the partition of unity: sum_i L_i(tau) = 1, so the points add up to the generator
    one = cref.ints_to_mont([1] * n, B.R)
    assert np.array_equal(cref.g1_jac_to_aff(ctx.msm_g1(one, got)), gen)


def test_rejects_bad_generator(ctx, pkg, cref):
    n = 8
    _, srs = _srs(ctx, cref, n, 5)
    with pytest.raises(pkg.B200Error):
        ctx.kzg_to_lagrange_g1(srs, cref.fr(N.Domain(16).generator))     # order 16, not 8
    with pytest.raises(pkg.B200Error):
        ctx.kzg_to_lagrange_g1(srs[:6], cref.fr(N.Domain(8).generator))  # not a power of two
