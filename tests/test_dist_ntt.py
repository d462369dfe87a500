"""Distributed computeH (one all-to-all per transform) without a GPU: the orchestration of
gnark-fork_b200/sharding.py over (a) virtual ranks in threads and (b) two gloo processes, the device
phases replayed by tests/cpu_model with the kernels' own index maps, against oracle computeH."""
import os
import socket
import subprocess
import sys
import textwrap
import threading

import numpy as np
import pytest

from oracle import ntt as N
from tests.conftest import ROOT, rand_field
from tests.dist_model import ModelEngine


class ThreadExchange:
    """all-to-all between `world` threads: chunk q of rank r's send buffer -> slice r of rank q's recv buffer."""

    def __init__(self, world):
        self.world, self.barrier, self.slots = world, threading.Barrier(world), [None] * world

    def bind(self, rank):
        def exchange(send, recv):
            self.slots[rank] = send
            self.barrier.wait()
            chunk = send.shape[0] // self.world
            for r in range(self.world):
                recv[r * chunk:(r + 1) * chunk] = self.slots[r][rank * chunk:(rank + 1) * chunk]
            self.barrier.wait()
        return exchange


@pytest.mark.parametrize("log_n,world,tl,smax", [(4, 2, 4, 3), (6, 2, 3, 2), (6, 4, 4, 3), (6, 8, 3, 3), (8, 4, 4, 2), (8, 2, 4, 3),
                                                 (9, 8, 4, 3), (10, 2, 9, 8), (10, 8, 5, 2)])
def test_distributed_compute_h_virtual_ranks(pkg, model, cref, log_n, world, tl, smax):
    n, n_loc = 1 << log_n, (1 << log_n) // world
    rng = np.random.default_rng(40 + log_n + world)
    a, b, c = (rand_field(rng, n) for _ in range(3))
    a[n - 3:] = 0; b[n - 3:] = 0; c[n - 3:] = 0           # zero padding like nbConstraints < n
    d = N.Domain(n)
    want = cref.compute_h(a, b, c, log_n, cref.fr(d.generator), cref.fr(d.coset_gen))
    ex = ThreadExchange(world)
    got = [None] * world
    errs = []

    def run(rank):
        try:
            eng = ModelEngine(model, cref, log_n, rank, world, tl, smax)
            blk = [v[rank * n_loc:(rank + 1) * n_loc].copy() for v in (a, b, c)]
            tmp = np.zeros((n_loc, 4), dtype=np.uint64)
            if (log_n + world) % 2:                       # both scratch arrangements
                tmp = [tmp, np.zeros_like(tmp), np.zeros_like(tmp)]
            got[rank] = pkg.sharding.compute_h_distributed(eng, ex.bind(rank), blk[0], blk[1], blk[2], tmp)
        except Exception as e:   # noqa: BLE001
            errs.append(e)
            ex.barrier.abort()

    ts = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    assert np.array_equal(np.concatenate(got), want)


def test_z_shard_range_is_block_aligned(pkg):
    sh = pkg.sharding
    assert sh.z_shard_range(15, 4, 0, 1) == (0, 15)
    assert [sh.z_shard_range(15, 4, i, 4) for i in range(4)] == [(0, 4), (4, 8), (8, 12), (12, 15)]
    assert [sh.z_shard_range(16, 4, i, 2) for i in range(2)] == [(0, 8), (8, 16)]
    assert [sh.z_shard_range(15, 4, i, 3) for i in range(3)] == [sh.shard_range(15, i, 3) for i in range(3)]
    with pytest.raises(ValueError):
        sh.log2_ranks(3)


WORKER = textwrap.dedent("""
    import os, sys, ctypes
    import numpy as np
    import torch
    import torch.distributed as dist
    sys.path.insert(0, {root!r})
    from tests.conftest import load_pkg, rand_field
    from tests.dist_model import ModelEngine
    from oracle import cref, ntt as N
    pkg = load_pkg()
    model = ctypes.CDLL({model!r})
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    log_n = 8
    n, n_loc = 1 << log_n, (1 << log_n) // world
    rng = np.random.default_rng(77)                      # same data on every rank
    a, b, c = (rand_field(rng, n) for _ in range(3))
    eng = ModelEngine(model, cref, log_n, rank, world, 4, 3)
    bufs = [torch.from_numpy(v[rank * n_loc:(rank + 1) * n_loc].copy().view(np.int64)) for v in (a, b, c)]
    tmp = [torch.zeros((n_loc, 4), dtype=torch.int64) for _ in range(3)]
    h = pkg.sharding.compute_h_distributed(eng, pkg.sharding.torch_exchange(dist, async_op=True), bufs[0], bufs[1], bufs[2], tmp)
    d = N.Domain(n)
    want = cref.compute_h(a, b, c, log_n, cref.fr(d.generator), cref.fr(d.coset_gen))
    lo, hi = rank * n_loc, (rank + 1) * n_loc
    assert np.array_equal(h.numpy().view(np.uint64), want[lo:hi]), "h block differs from the oracle"
    assert pkg.sharding.z_shard_range(n - 1, log_n, rank, world) == (lo, min(hi, n - 1))
    dist.barrier()
    dist.destroy_process_group()
    print("rank", rank, "ok")
""")


def test_two_rank_gloo_distributed_compute_h(tmp_path, model):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT, model=os.path.join(ROOT, "tests", "cpu_model", "libmodel.so")))
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), WORLD_SIZE="2", OMP_NUM_THREADS="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r), LOCAL_RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r} failed:\n{o}"
        assert f"rank {r} ok" in o
