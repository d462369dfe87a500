"""A few transforms of one size (development aid for ncu): python scripts/one_ntt.py <log_n>"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, fr_limbs, FR_ROOT_2_28, R_MOD
pkg = load_pkg()
ctx = pkg.Context(0)
log_n = int(sys.argv[1])
n = 1 << log_n
gen = fr_limbs(pow(FR_ROOT_2_28, 1 << (28 - log_n), R_MOD)); cg = fr_limbs(5)
v = pkg.DeviceVector(ctx, n * 32).upload(rand_field(np.random.default_rng(0), n))
dif = [ctx.ntt_dev(v, log_n, 0, 0, 0, gen, cg) for _ in range(3)]          # forward DIF, natural -> bit-reversed
dit = [ctx.ntt_dev(v, log_n, 1, 1, 1, gen, cg) for _ in range(3)]          # inverse DIT on the coset
print("ntt 2^%d dif %.3f ms, inverse coset dit %.3f ms" % (log_n, min(dif[1:]), min(dit[1:])))
