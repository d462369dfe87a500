"""A/B of the G2 bucket accumulation: registers (default) vs running sum in shared memory (msm_acc_smem bit 1).
python scripts/g2_smem_ab.py [log_n=22] [curve]   (B200PROVER_LIB selects a library build)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import load_pkg, rand_field, BN254_PARAMS, BLS12_377_PARAMS
pkg = load_pkg()
log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 22
cp = BLS12_377_PARAMS if len(sys.argv) > 2 and sys.argv[2] == "bls12_377" else BN254_PARAMS
ctx = pkg.Context(0, curve=cp.id)
rng = np.random.default_rng(0)
n = 1 << log_n
bh = ctx.upload_bases(ctx.batch_scalar_mul(cp.g2_gen, rand_field(rng, n, top_limb=cp.top_limb), 2), 2)
sc = rand_field(rng, n, top_limb=cp.top_limb)
sc[:100] = sc[0]          # repeated scalars
sv = pkg.DeviceVector(ctx, n * 32).upload(sc)
res = {}
for mode in (0, 1, 0, 1):
    ctx.set_option("msm_acc_smem", 2 * mode)
    ctx.profile_start()
    for _ in range(3):
        out, ms = ctx.msm_dev(bh, sv, n)
    p = ctx.profile_read()["msm_acc_g2"]
    res[mode] = pkg.g2_jac_to_affine(out, curve=cp.id)
    print(f"{cp.name} G2 2^{log_n} smem={mode}: msm {ms:.2f} ms, accumulate kernel {p['ms'] / max(1, p['launches']):.2f} ms")
print("results equal:", np.array_equal(res[0], res[1]))
