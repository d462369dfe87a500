"""Folded-window sweep with per-phase times (development aid): python scripts/fold_sweep.py 20,22 [group]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, G1_GEN, G2_GEN
pkg = load_pkg()
ctx = pkg.Context(0)
rng = np.random.default_rng(0)
group = int(sys.argv[2]) if len(sys.argv) > 2 else 1
for log_n in [int(x) for x in sys.argv[1].split(",")]:
    n = 1 << log_n
    bases = ctx.batch_scalar_mul(G1_GEN if group == 1 else G2_GEN, rand_field(rng, n), group)
    sv = pkg.DeviceVector(ctx, n * 32).upload(rand_field(rng, n))
    for c in [0] + list(range(max(10, log_n - 6), min(24, log_n + 1))):
        ctx.set_option("msm_precompute", 0 if c == 0 else 1)
        ctx.set_option("msm_fold_c", c)
        try:
            bh = ctx.upload_bases(bases, group)
        except pkg.B200Error as e:
            print(log_n, c, "ERR", e); continue
        ctx.msm_dev(bh, sv, n)
        ctx.profile_start()
        ms = min(ctx.msm_dev(bh, sv, n)[1] for _ in range(2))
        pr = ctx.profile_read()
        acc = pr["msm_acc_g1"]["ms"] + pr["msm_acc_g2"]["ms"]
        print(f"g{group} 2^{log_n} fold_c={c:2d} total {ms:7.3f} ms | sort {pr['msm_sort']['ms']/2:6.3f} acc {acc/2:7.3f} reduce {pr['msm_reduce']['ms']/2:6.3f}", flush=True)
        bh.free()
    sv.free()
