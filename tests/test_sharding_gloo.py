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
