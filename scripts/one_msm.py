"""One MSM (development aid for ncu): python scripts/one_msm.py <group 1|2> <log_n> [batched-affine rounds] [curve: bn254|bls12_377]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, BN254_PARAMS, BLS12_377_PARAMS
pkg = load_pkg()
group, log_n = int(sys.argv[1]), int(sys.argv[2])
cp = BLS12_377_PARAMS if len(sys.argv) > 4 and sys.argv[4] == "bls12_377" else BN254_PARAMS
G1_GEN, G2_GEN = cp.g1_gen, cp.g2_gen
ctx = pkg.Context(0, curve=cp.id)
if len(sys.argv) > 3 and int(sys.argv[3]):
    ctx.set_option("msm_ba", int(sys.argv[3]))
for kv in filter(None, os.environ.get("B200_OPTS", "").split(",")):     # e.g. B200_OPTS=msm_acc_smem=3
    ctx.set_option(kv.split("=")[0], int(kv.split("=")[1]))
rng = np.random.default_rng(0)
n = 1 << log_n
bh = ctx.upload_bases(ctx.batch_scalar_mul(G1_GEN if group == 1 else G2_GEN, rand_field(rng, n, top_limb=cp.top_limb), group), group)
sv = pkg.DeviceVector(ctx, n * 32).upload(rand_field(rng, n, top_limb=cp.top_limb))
for _ in range(2):
    out, ms = ctx.msm_dev(bh, sv, n)
print("msm group", group, "2^%d" % log_n, ms, "ms")
