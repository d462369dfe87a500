"""Device-side timing probe (development aid, not the bench): IMAD peak, NTT and MSM sweeps with
CUDA-event times from the C ABI's *_dev entry points.  Usage: python scripts/gpu_probe.py [max_log]"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests.conftest import load_pkg, rand_field  # noqa: E402
from oracle import bn254 as B  # noqa: E402
from oracle import ntt as N  # noqa: E402
from oracle import cref  # noqa: E402

pkg = load_pkg()
max_log = int(sys.argv[1]) if len(sys.argv) > 1 else 22
ctx = pkg.Context(0)
res = {"peak": ctx.imad_peak()}
print(json.dumps(res["peak"]), flush=True)
rng = np.random.default_rng(0)
gen1 = cref.g1_points_to_array([B.G1_GEN])[0]
gen2 = cref.g2_points_to_array([B.G2_GEN])[0]
res["ntt"], res["msm_g1"], res["msm_g2"] = {}, {}, {}
for log_n in range(16, max_log + 1, 2):
    n = 1 << log_n
    d = N.Domain(n)
    gen, cg = cref.fr(d.generator), cref.fr(d.coset_gen)
    x = rand_field(rng, n)
    v = pkg.DeviceVector(ctx, n * 32).upload(x)
    ts = [ctx.ntt_dev(v, log_n, 0, 0, 0, gen, cg) for _ in range(4)]
    tc = [ctx.ntt_dev(v, log_n, 1, 0, 1, gen, cg) for _ in range(3)]
    res["ntt"][log_n] = {"fwd_dif_ms": min(ts[1:]), "inv_dif_coset_ms": min(tc[1:])}
    print("ntt", log_n, res["ntt"][log_n], flush=True)
    v.free()
    t0 = time.time()
    bases = ctx.batch_scalar_mul(gen1, x, 1)
    t_bm = time.time() - t0
    bh = ctx.upload_bases(bases, 1)
    sv = pkg.DeviceVector(ctx, n * 32).upload(rand_field(rng, n))
    ts = [ctx.msm_dev(bh, sv, n)[1] for _ in range(3)]
    res["msm_g1"][log_n] = {"ms": min(ts[1:]), "points_per_s": n / (min(ts[1:]) * 1e-3), "batch_mul_s": t_bm}
    print("msm_g1", log_n, res["msm_g1"][log_n], flush=True)
    bh.free()
    if log_n <= 20:
        bases2 = ctx.batch_scalar_mul(gen2, x, 2)
        bh2 = ctx.upload_bases(bases2, 2)
        ts = [ctx.msm_dev(bh2, sv, n)[1] for _ in range(3)]
        res["msm_g2"][log_n] = {"ms": min(ts[1:]), "points_per_s": n / (min(ts[1:]) * 1e-3)}
        print("msm_g2", log_n, res["msm_g2"][log_n], flush=True)
        bh2.free()
    sv.free()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/probe.json", "w"), indent=1)
ctx.close()
