"""Batched-affine rounds sweep: device time of one resident-base MSM for msm_ba = 0..5 (results must agree).
python scripts/ba_sweep.py <group 1|2> <log_n> [rounds, comma separated]"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, G1_GEN, G2_GEN
pkg = load_pkg()
group, log_n = int(sys.argv[1]), int(sys.argv[2])
ctx = pkg.Context(0)
rng = np.random.default_rng(0)
n = 1 << log_n
bh = ctx.upload_bases(ctx.batch_scalar_mul(G1_GEN if group == 1 else G2_GEN, rand_field(rng, n), group), group)
sv = pkg.DeviceVector(ctx, n * 32).upload(rand_field(rng, n))
aff = pkg.g1_jac_to_affine if group == 1 else pkg.g2_jac_to_affine
ref, res = None, {}
rounds = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else list(range(0, 6))
for ba in rounds:
    ctx.set_option("msm_ba", ba)
    ts = []
    for _ in range(4):
        out, ms = ctx.msm_dev(bh, sv, n)
        ts.append(ms)
    ctx.profile_start()
    out, ms = ctx.msm_dev(bh, sv, n)
    prof = ctx.profile_read()
    a = aff(out)
    if ref is None:
        ref = a
    res[ba] = {"ms": min(ts[1:]), "same_result": bool(np.array_equal(a, ref)),
               "acc_ms": prof["msm_acc_g1" if group == 1 else "msm_acc_g2"]["ms"], "sort_ms": prof["msm_sort"]["ms"], "reduce_ms": prof["msm_reduce"]["ms"]}
    print(ba, res[ba], flush=True)
print(json.dumps({"group": group, "log_n": log_n, "by_rounds": res}))
