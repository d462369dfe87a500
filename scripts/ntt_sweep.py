"""NTT launch-shape sweep (development aid)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, fr_limbs, FR_ROOT_2_28, R_MOD
pkg = load_pkg()
ctx = pkg.Context(0)
rng = np.random.default_rng(0)
for log_n in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "20,24").split(",")]:
    n = 1 << log_n
    gen = fr_limbs(pow(FR_ROOT_2_28, 1 << (28 - log_n), R_MOD)); cg = fr_limbs(5)
    v = pkg.DeviceVector(ctx, n * 32).upload(rand_field(rng, n))
    for tl in (8, 9, 10, 11, 12):
        for th in (128, 256, 512, 1024):
            for sm in (6, 8, 12):
                if sm > tl: continue
                ctx.set_option("ntt_tile_log", tl); ctx.set_option("ntt_strided_stages", sm); ctx.set_option("ntt_threads", th)
                ts = [ctx.ntt_dev(v, log_n, 0, 0, 0, gen, cg) for _ in range(3)]
                td = [ctx.ntt_dev(v, log_n, 1, 1, 1, gen, cg) for _ in range(3)]
                print("ntt", log_n, "tile", tl, "threads", th, "strided", sm, "dif", round(min(ts[1:]), 3), "dit_inv_coset", round(min(td[1:]), 3), flush=True)
    v.free()
