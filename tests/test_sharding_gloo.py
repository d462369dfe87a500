"""N > 1 host logic on the CPU: world_size-2 gloo processes, MSM sharding by point range,
all_gather of the partial sums and host combine (gnark-fork_b200/sharding.py).  The per-shard MSM is
computed by the oracle here (no GPU in this test): what is under test is the shard split, the
gather and b200_g1/g2_jac_add."""
import os
import socket
import subprocess
import sys
import textwrap

from tests.conftest import ROOT

WORKER = textwrap.dedent("""
    import os, sys
    import numpy as np
    import torch.distributed as dist
    sys.path.insert(0, {root!r})
    from tests.conftest import load_pkg, rand_field
    from tests import util
    from oracle import cref
    pkg = load_pkg()
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    n = 301
    rng = np.random.default_rng(5)                      # same data on every rank
    scalars = rand_field(rng, n)
    b1, _ = util.known_dlog_bases_g1(n, seed=2)
    b2, _ = util.known_dlog_bases_g2(n, seed=2)
    lo, hi = pkg.sharding.shard_range(n, rank, world)
    part = dict(ar=cref.msm_g1(scalars[lo:hi], b1[lo:hi]), bs1=cref.msm_g1(scalars[lo:hi], b1[lo:hi]),
                krs=np.zeros(12, np.uint64), krs2=cref.msm_g1(scalars[lo:hi], b1[lo:hi]),
                bs2=cref.msm_g2(scalars[lo:hi], b2[lo:hi]))
    comb = pkg.sharding.gather_combine(dist, part)
    full1, full2 = cref.msm_g1(scalars, b1), cref.msm_g2(scalars, b2)
    assert np.array_equal(pkg.g1_jac_to_affine(comb["ar"]), cref.g1_jac_to_aff(full1))
    assert np.array_equal(pkg.g1_jac_to_affine(comb["krs2"]), cref.g1_jac_to_aff(full1))
    assert not pkg.g1_jac_to_affine(comb["krs"]).any()            # infinity + infinity
    assert np.array_equal(pkg.g2_jac_to_affine(comb["bs2"]), cref.g2_jac_to_aff(full2))
    ranges = [pkg.sharding.shard_range(n, i, world) for i in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == n and all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
    dist.barrier()
    dist.destroy_process_group()
    print("rank", rank, "ok")
""")


def test_two_rank_partial_sum_combine(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT))
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), WORLD_SIZE="2", OMP_NUM_THREADS="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r), LOCAL_RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r} failed:\n{o}"
        assert f"rank {r} ok" in o


PLONK_WORKER = textwrap.dedent("""
    import os, sys, random
    import numpy as np
    import torch.distributed as dist
    sys.path.insert(0, {root!r})
    from tests.conftest import load_pkg
    from tests import util
    from oracle import plonk_prove as PP, ntt as N
    from oracle.groth16 import g1_gen_mul
    pkg = load_pkg()
    S = pkg.sharding
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    R = PP.R
    rng = random.Random(11)                              # same data on every rank
    tau, m = rng.randrange(R), 37
    polys = [[rng.randrange(R) for _ in range(m)] for _ in range(3)]
    lo, hi = m * rank // world, m * (rank + 1) // world  # the SRS point range of this rank (plonk.cu: srs_lo_idx / srs_hi_idx)
    def jac(p):                                          # affine ints -> G1Jac limbs with Z = 1
        a = util.g1_arr([p])[0]
        one = util.g1_arr([(1, 2)])[0][:4]               # Montgomery form of 1 is the X limb of the generator (1, 2)
        return np.concatenate([a, one]) if p is not None else np.zeros(12, np.uint64)
    partial = np.stack([jac(g1_gen_mul(sum(c * pow(tau, i, R) for i, c in enumerate(p) if lo <= i < hi) % R)) for p in polys])
    total = S._gather_points(dist, partial)
    for j, p in enumerate(polys):
        want = PP.commit_tau(p, tau)
        assert util.g1_aff_ints(pkg.g1_jac_to_affine(total[j])) == list(want)
    # coset k of the numerator lands in block bitrev(k): prove.go:1029-1035 writes cres[bitrev(rho * i + k)]
    for log_n, lb in ((3, 2), (4, 3)):
        n, rho = 1 << log_n, 1 << lb
        for k in range(rho):
            for i in (0, 1, n - 1):
                pos = N.bitrev(rho * i + k, log_n + lb)
                assert pos // n == S.coset_block(k, lb) and pos % n == N.bitrev(i, log_n)
    dist.barrier()
    dist.destroy_process_group()
    print("rank", rank, "ok")
""")


def test_two_rank_plonk_partial_commitments(tmp_path):
    """Host logic of the sharded PlonK proof (sharding.py): partial commitments over SRS point ranges gathered and
    summed == the full commitment; the coset-block map of the numerator exchange."""
    script = tmp_path / "worker_plonk.py"
    script.write_text(PLONK_WORKER.format(root=ROOT))
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), WORLD_SIZE="2", OMP_NUM_THREADS="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r), LOCAL_RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r} failed:\n{o}"
        assert f"rank {r} ok" in o
