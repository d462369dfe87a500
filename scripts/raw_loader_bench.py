"""Time the on-disk -> device conversion of a point slice (b200_bases_upload_raw) against the in-memory upload.

    python scripts/raw_loader_bench.py [log_n=22] [curve=bn254|bls12_377]

The raw slice is what ProvingKey.WriteRawTo stores for pk.G1.A (big-endian canonical X || Y); it is made here from
device-generated points (canonical limbs = Montgomery limbs x 1, byte-reversed per coordinate).  Folded-window tables
are switched off so that the two uploads differ by the conversion kernel only; the converted bases are checked by an
MSM against the in-memory ones."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, BN254_PARAMS, BLS12_377_PARAMS
pkg = load_pkg()
log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 22
cp = BLS12_377_PARAMS if len(sys.argv) > 2 and sys.argv[2] == "bls12_377" else BN254_PARAMS
ctx = pkg.Context(0, curve=cp.id)
ctx.set_option("msm_precompute", 0)
fw, n = ctx.fw, 1 << log_n
rng = np.random.default_rng(0)
for group in (1, 2):
    m = n if group == 1 else n // 2
    bases = ctx.batch_scalar_mul(cp.g1_gen if group == 1 else cp.g2_gen, rand_field(rng, m, top_limb=cp.top_limb), group)
    coords = bases.reshape(-1, fw)
    one = np.zeros((coords.shape[0], fw), dtype=np.uint64); one[:, 0] = 1
    canon = ctx.fp_vec_mul(coords, one)                              # x R * 1 / R = x: canonical little-endian limbs
    be = canon.view(np.uint8).reshape(-1, 8 * fw)[:, ::-1]           # big-endian bytes per coordinate
    k = 2 if group == 1 else 4
    be = be.reshape(m, k, 8 * fw)
    if group == 2:
        be = be[:, [1, 0, 3, 2], :]                                  # disk order X.A1 X.A0 Y.A1 Y.A0
    raw = pkg.pinned_empty((m * k * 8 * fw,), dtype=np.uint8)
    raw[:] = np.ascontiguousarray(be).reshape(-1)
    pin = pkg.pinned_empty(bases.shape)
    pin[:] = bases
    res = {}
    for name, fn in (("in-memory limbs", lambda: ctx.upload_bases(pin, group)), ("raw file bytes", lambda: ctx.upload_bases_raw(raw, m, group))):
        fn().free()
        ts = []
        for _ in range(5):
            ctx.sync(); t0 = time.perf_counter(); h = fn(); ctx.sync(); ts.append(time.perf_counter() - t0); last = h
            if _ < 4: h.free()
        res[name] = (min(ts), last)
    sc = rand_field(rng, m, top_limb=cp.top_limb)
    a = ctx.msm_resident(res["in-memory limbs"][1], sc)
    b = ctx.msm_resident(res["raw file bytes"][1], sc)
    aff = pkg.g1_jac_to_affine if group == 1 else pkg.g2_jac_to_affine
    ok = np.array_equal(aff(a, curve=cp.id), aff(b, curve=cp.id))
    gb = raw.nbytes / 1e9
    print(f"{cp.name} G{group} 2^{int(np.log2(m))} points ({gb:.3f} GB): in-memory {res['in-memory limbs'][0]*1e3:.2f} ms, "
          f"raw {res['raw file bytes'][0]*1e3:.2f} ms (+{(res['raw file bytes'][0]-res['in-memory limbs'][0])*1e3:.2f} ms conversion), MSM equal: {ok}")
    for v in res.values():
        v[1].free()
