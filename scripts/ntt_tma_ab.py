"""A/B of the TMA-staged NTT pass against the per-thread-load pass (development aid; numbers quoted in DESIGN.md):
python scripts/ntt_tma_ab.py [log_n,...]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, fr_limbs, FR_ROOT_2_28, R_MOD
pkg = load_pkg()
ctx = pkg.Context(0)
rng = np.random.default_rng(0)
for log_n in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "20,22,24,26").split(",")]:
    n = 1 << log_n
    gen = fr_limbs(pow(FR_ROOT_2_28, 1 << (28 - log_n), R_MOD)); cg = fr_limbs(5)
    v = pkg.DeviceVector(ctx, n * 32).upload(rand_field(rng, n))
    for tma in (0, 1):
        for tl, th, sm in ((9, 128, 8), (9, 256, 8), (10, 256, 8), (10, 512, 8), (11, 512, 8), (11, 1024, 8), (12, 1024, 8), (12, 1024, 6)):
            ctx.set_option("ntt_tma", tma); ctx.set_option("ntt_tile_log", tl); ctx.set_option("ntt_strided_stages", sm); ctx.set_option("ntt_threads", th)
            ts = [ctx.ntt_dev(v, log_n, 1, 0, 0, gen, cg) for _ in range(4)]          # FFTInverse DIF (prove.go:369)
            td = [ctx.ntt_dev(v, log_n, 0, 1, 1, gen, cg) for _ in range(4)]          # FFT DIT on coset (prove.go:373)
            print("ntt 2^%d tma=%d tile=%d threads=%d strided=%d  inv_dif %.3f ms  fwd_dit_coset %.3f ms" % (log_n, tma, tl, th, sm, min(ts[1:]), min(td[1:])), flush=True)
    v.free()
