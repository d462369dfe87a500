"""Tuning sweep on the GPU box: MSM window width and NTT tile plan (development aid)."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, fr_limbs, G1_GEN, G2_GEN, FR_ROOT_2_28, R_MOD  # noqa: E402

pkg = load_pkg()
ctx = pkg.Context(0)
rng = np.random.default_rng(0)
res = {"msm_g1": {}, "msm_g2": {}, "ntt": {}}
logs = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "20,22,24").split(",")]
for log_n in logs:
    n = 1 << log_n
    k = rand_field(rng, n)
    bh = ctx.upload_bases(ctx.batch_scalar_mul(G1_GEN, k, 1), 1)
    sv = pkg.DeviceVector(ctx, n * 32).upload(rand_field(rng, n))
    for c in range(13, 23):
        ctx.set_option("msm_c", c)
        try:
            ts = [ctx.msm_dev(bh, sv, n)[1] for _ in range(3)]
        except pkg.B200Error as e:
            print("g1", log_n, c, "ERR", e)
            continue
        res["msm_g1"][f"{log_n},{c}"] = min(ts[1:])
        print("msm_g1", log_n, c, round(min(ts[1:]), 3), flush=True)
    bh.free()
    if log_n <= 22:
        bh2 = ctx.upload_bases(ctx.batch_scalar_mul(G2_GEN, k, 2), 2)
        for c in range(13, 21):
            ctx.set_option("msm_c", c)
            ts = [ctx.msm_dev(bh2, sv, n)[1] for _ in range(2)]
            res["msm_g2"][f"{log_n},{c}"] = min(ts[1:])
            print("msm_g2", log_n, c, round(min(ts[1:]), 3), flush=True)
        bh2.free()
    sv.free()
    ctx.set_option("msm_c", 0)
    gen = fr_limbs(pow(FR_ROOT_2_28, 1 << (28 - log_n), R_MOD))
    cg = fr_limbs(5)
    v = pkg.DeviceVector(ctx, n * 32).upload(rand_field(rng, n))
    for tl in (10, 11, 12):
        for sm in (4, 6, 8, 10, 12):
            if sm > tl:
                continue
            ctx.set_option("ntt_tile_log", tl)
            ctx.set_option("ntt_strided_stages", sm)
            ts = [ctx.ntt_dev(v, log_n, 0, 0, 0, gen, cg) for _ in range(3)]
            td = [ctx.ntt_dev(v, log_n, 0, 1, 0, gen, cg) for _ in range(3)]
            res["ntt"][f"{log_n},{tl},{sm}"] = [min(ts[1:]), min(td[1:])]
            print("ntt", log_n, "tile", tl, "strided", sm, "dif", round(min(ts[1:]), 3), "dit", round(min(td[1:]), 3), flush=True)
    v.free()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/tune.json", "w"), indent=1)
