"""One MSM (development aid for ncu): python scripts/one_msm.py <group 1|2> <log_n> [batched-affine rounds]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, G1_GEN, G2_GEN
pkg = load_pkg()
group, log_n = int(sys.argv[1]), int(sys.argv[2])
ctx = pkg.Context(0)
if len(sys.argv) > 3:
    ctx.set_option("msm_ba", int(sys.argv[3]))
rng = np.random.default_rng(0)
n = 1 << log_n
bh = ctx.upload_bases(ctx.batch_scalar_mul(G1_GEN if group == 1 else G2_GEN, rand_field(rng, n), group), group)
sv = pkg.DeviceVector(ctx, n * 32).upload(rand_field(rng, n))
for _ in range(2):
    out, ms = ctx.msm_dev(bh, sv, n)
print("msm group", group, "2^%d" % log_n, ms, "ms")
