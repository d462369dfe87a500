"""Shared helpers for the parity tests: golden fixture access and limb conversions."""
import json
import os

import numpy as np

from oracle import bn254 as B
from oracle import cref

_G = None


def golden():
    global _G
    if _G is None:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bn254_small.json")) as f:
            _G = json.load(f)
    return _G


def fr_arr(xs):
    return cref.ints_to_mont([x % B.R for x in xs], B.R)


def fr_ints(a):
    return cref.mont_to_ints(np.asarray(a), B.R)


def g1_arr(pts):
    return cref.g1_points_to_array([None if p is None else (p[0], p[1]) for p in pts])


def g2_arr(pts):
    return cref.g2_points_to_array([None if p is None else ((p[0], p[1]), (p[2], p[3])) for p in pts])


def g1_aff_ints(aff8):
    p = cref.g1_array_to_points(np.asarray(aff8).reshape(1, 8))[0]
    return None if p is None else [p[0], p[1]]


def g2_aff_ints(aff16):
    p = cref.g2_array_to_points(np.asarray(aff16).reshape(1, 16))[0]
    return None if p is None else [p[0][0], p[0][1], p[1][0], p[1][1]]


def known_dlog_bases_g1(n, seed=1):
    """n distinct G1 points with known discrete logs (oracle fixed-base batch mul)."""
    rng = np.random.default_rng(seed)
    ks = [int(x) for x in rng.integers(1, 1 << 62, size=n)]
    gen = cref.g1_points_to_array([B.G1_GEN])[0]
    return cref.g1_batch_mul(gen, cref.ints_to_mont(ks, B.R)), ks


def known_dlog_bases_g2(n, seed=1):
    rng = np.random.default_rng(seed)
    ks = [int(x) for x in rng.integers(1, 1 << 62, size=n)]
    gen = cref.g2_points_to_array([B.G2_GEN])[0]
    return cref.g2_batch_mul(gen, cref.ints_to_mont(ks, B.R)), ks
