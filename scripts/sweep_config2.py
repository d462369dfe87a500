"""BASELINE.json config 2: standalone G1 MSM sweep 2^16..2^26 (resident bases, uniform and witness-like
scalars) and Fr NTT / coset / inverse sweep 2^16..2^26, device times from CUDA events.
    python scripts/sweep_config2.py [max_log] > gpurun_out/sweep_config2.json
Results are checked: MSM against the known discrete logs of the bases, NTT by a forward/inverse round trip."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, fr_limbs, mont_sum, G1_GEN, FR_ROOT_2_28, R_MOD  # noqa: E402

pkg = load_pkg()
ctx = pkg.Context(0)
rng = np.random.default_rng(26)
max_log = int(sys.argv[1]) if len(sys.argv) > 1 else 26
peak = ctx.imad_peak()
out = {"imad_peak": peak, "msm_g1": {}, "ntt": {}}


def witness_like(n):
    """50 % in {0, 1}, 25 % below 2^16, 25 % uniform (Montgomery form)."""
    s = rand_field(rng, n)
    kind = rng.integers(0, 4, size=n)
    small = np.zeros((n, 4), dtype=np.uint64)
    small[:, 0] = np.where(kind < 2, rng.integers(0, 2, size=n), rng.integers(0, 1 << 16, size=n)).astype(np.uint64)
    r2 = np.tile(fr_limbs(pow(2, 512, R_MOD)), (n, 1))          # to Montgomery form: x * R^2 * R^-1
    small_m = ctx.fr_vec_mul(small, r2)
    s[kind < 3] = small_m[kind < 3]
    return s


for log_n in range(16, max_log + 1, 2):
    n = 1 << log_n
    k = rand_field(rng, n)
    bases = ctx.batch_scalar_mul(G1_GEN, k, 1)
    bh = ctx.upload_bases(bases, 1)
    info = bh.info()
    del bases
    row = {"plan": info}
    for name, scal in (("uniform", rand_field(rng, n)), ("witness_like", witness_like(n))):
        sv = pkg.DeviceVector(ctx, n * 32).upload(scal)
        res, _ = ctx.msm_dev(bh, sv, n)
        ms = min(ctx.msm_dev(bh, sv, n)[1] for _ in range(3))
        tot = mont_sum(ctx.fr_vec_mul(k, scal))
        want = ctx.batch_scalar_mul(G1_GEN, np.frombuffer(tot.to_bytes(32, "little"), dtype=np.uint64).reshape(1, 4), 1)[0]
        ok = bool(np.array_equal(pkg.g1_jac_to_affine(res), want))
        row[name] = {"ms": ms, "points_per_s": n / (ms * 1e-3), "known_dlog_ok": ok,
                     "frac_imad_survey_model": n * 23936 / (ms * 1e-3) / peak["mac32_per_s"]}
        sv.free()
        print(f"msm_g1 2^{log_n} {name}: {ms:.3f} ms ok={ok}", file=sys.stderr, flush=True)
    out["msm_g1"][str(log_n)] = row
    bh.free()
    del k

    gen = fr_limbs(pow(FR_ROOT_2_28, 1 << (28 - log_n), R_MOD))
    cg = fr_limbs(5)
    x = rand_field(rng, n)
    v = pkg.DeviceVector(ctx, n * 32).upload(x)
    r = {}
    for name, (inv, dec, coset) in {"fft_dif": (0, 0, 0), "fft_dit_coset": (0, 1, 1), "ifft_dif": (1, 0, 0), "ifft_dif_coset": (1, 0, 1)}.items():
        ts = [ctx.ntt_dev(v, log_n, inv, dec, coset, gen, cg) for _ in range(4)]
        ms = min(ts[1:])
        passes = 1 if log_n <= 12 else (2 if log_n <= 24 else 3)
        r[name] = {"ms": ms, "hbm_gbs_survey_model": n * 64.0 * passes / (ms * 1e-3) / 1e9,
                   "frac_imad": n * 0.5 * log_n * 136 / (ms * 1e-3) / peak["mac32_per_s"]}
    # round trip on fresh data: inverse DIF (natural -> bit-reversed) then forward DIT (bit-reversed -> natural)
    v.upload(x)
    ctx.ntt_dev(v, log_n, 1, 0, 0, gen, cg)
    ctx.ntt_dev(v, log_n, 0, 1, 0, gen, cg)
    r["round_trip_ok"] = bool(np.array_equal(v.download((n, 4)), x))
    out["ntt"][str(log_n)] = r
    print(f"ntt 2^{log_n}: {r['fft_dif']['ms']:.3f} ms round trip {r['round_trip_ok']}", file=sys.stderr, flush=True)
    v.free()
print(json.dumps(out))
