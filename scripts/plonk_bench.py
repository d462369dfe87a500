"""PlonK building blocks at BASELINE.json config 4 sizes (n = 2^22 gates, big domain 2^24): wall time of
each synchronous ABI call after warm-up, bytes model vs the measured HBM peak, and the quotient stage
(P4 + P5 + P6) assembled from them on synthetic data.   python scripts/plonk_bench.py [log_n] > out.json"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, fr_limbs, G1_GEN, FR_ROOT_2_28, R_MOD, ROOT  # noqa: E402

pkg = load_pkg()
ctx = pkg.Context(0)
log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 22
log_big = log_n + 2
n, big = 1 << log_n, 1 << log_big
rng = np.random.default_rng(4)
peaks = os.path.join(ROOT, "MEASURED_PEAKS.json")
hbm = float(json.load(open(peaks))["hbm_gbs"]) if os.path.exists(peaks) else 6650.0
imad = ctx.imad_peak()["mac32_per_s"]
gen = fr_limbs(pow(FR_ROOT_2_28, 1 << (28 - log_n), R_MOD))
gen_big = fr_limbs(pow(FR_ROOT_2_28, 1 << (28 - log_big), R_MOD))
cg = fr_limbs(5)
out = {"log_n": log_n, "log_big": log_big, "hbm_peak_gbs": hbm}


def timed(fn, reps=3):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts) * 1e3


def vec(m=n):
    return pkg.DeviceVector(ctx, m * 32).upload(rand_field(rng, m))


a, b = vec(), vec()
z = fr_limbs(0x123456789ABCDEF)
rows = {}
rows["batch_inverse"] = (timed(lambda: ctx.fr_batch_inverse_dev(a, b, n)), 128.0 * n)          # 2 reads + 2 writes of the scratch/out
rows["prefix_product"] = (timed(lambda: ctx.fr_prefix_product_dev(a, b, n)), 96.0 * n)         # 2 reads + 1 write
rows["poly_eval"] = (timed(lambda: ctx.fr_poly_eval_dev(a, n, z)), 32.0 * n)
rows["poly_div_linear"] = (timed(lambda: ctx.fr_poly_div_linear_dev(a, n, z, b)), 96.0 * n)
bigv = vec(big)
rows["divide_by_xn_minus_one (2^%d, scale + inverse coset NTT)" % log_big] = (
    timed(lambda: ctx.plonk_divide_by_xn_minus_one_dev(bigv, log_big, log_n, gen_big, cg)), 64.0 * big * 4)
perm = pkg.DeviceVector(ctx, 3 * n * 8).upload(rng.permutation(3 * n).astype(np.int64))
l, r, o = vec(), vec(), vec()
rows["build_ratio (Z)"] = (timed(lambda: ctx.plonk_build_ratio_dev(l, r, o, perm, z, z, log_n, gen, cg, b)), (3 * 32 + 24 + 7 * 32) * float(n))
x = [None if k == 4 else vec() for k in range(15)]
bl = rand_field(rng, 2); bz = rand_field(rng, 3)
t_rows = timed(lambda: ctx.plonk_quotient_rows_dev(x, [], log_n, log_big, 1, gen, z, z, z, cg, bl, bl, bl, bz, bigv))
rows["quotient_rows (one coset)"] = (t_rows, (15 * 32 + 32) * float(n))
out["quotient_rows_frac_imad"] = n * 26 * 136 / (t_rows * 1e-3) / imad
srs_k = rand_field(rng, n)
srs = ctx.upload_bases(ctx.batch_scalar_mul(G1_GEN, srs_k, 1), 1)
rows["kzg_commit (G1 MSM 2^%d, resident SRS)" % log_n] = (timed(lambda: ctx.kzg_commit_dev(srs, a, n)), 96.0 * n)
out["kernels"] = {k: {"ms": ms, "gbs_model": by / (ms * 1e-3) / 1e9, "frac_hbm": by / (ms * 1e-3) / 1e9 / hbm} for k, (ms, by) in rows.items()}

# quotient stage: 4 cosets x 14 polynomials x (inverse DIF + forward DIT on the coset) + rows, then the division
polys = [v for v in x if v is not None]


def quotient_stage():
    for coset in range(4):
        for p in polys:
            ctx.ntt_dev(p, log_n, 1, 0, 0, gen, cg)
            ctx.ntt_dev(p, log_n, 0, 1, 1, gen, cg)
        ctx.plonk_quotient_rows_dev(x, [], log_n, log_big, coset, gen, z, z, z, cg, bl, bl, bl, bz, bigv)
    ctx.plonk_divide_by_xn_minus_one_dev(bigv, log_big, log_n, gen_big, cg)


out["quotient_stage_ms"] = timed(quotient_stage, reps=2)
out["quotient_stage_note"] = "P4 + P5 + P6 on synthetic data: 112 transforms of 2^%d, 4 row passes, one 2^%d inverse coset transform" % (log_n, log_big)
print(json.dumps(out, indent=1))
