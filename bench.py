#!/usr/bin/env python
"""bench.py -- BN254 Groth16 prove core (computeH + the five MultiExp of
backend/groth16/bn254/prove.go:139-301) at 2^24 constraints on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--log-n 24] [--impl reference]

One JSON line on stdout (rank 0).  A "step" is one proof of the refCircuit instance (groth16_test.go:120-156: the
X <- X*X chain, 2^log_n constraints, its SATISFYING witness; SURVEY.md 8: every MSM has ~2^log_n points) with a key
made by b200_groth16_setup from a seeded toxic waste, so the proof is checked before timing: all five MSM results
against the discrete logs the witness implies, the verify equation in the exponent, the host fix-ups.  `value` =
seconds per proof with a, b, c, W already resident in HBM (CUDA events on the launching stream, max over ranks);
`e2e` = seconds per proof through the host-buffer C-ABI call the cgo shim makes (b200_groth16_prove_core from pinned
host memory + the host fix-ups), H2D/D2H inside the timed region.  N > 1 (torchrun, one process per GPU): strong
scaling -- the five MSMs are sharded by point range, partial sums gathered and added on the host; computeH is
distributed with one all-to-all per transform.
`cpu_baseline` (N = 1): one prove core of the C++ restatement (oracle/) at 2^20 on the same kind of instance, on all
host cores, compared with the device's results on the same inputs and scaled to 2^log_n.
`--impl reference`: no Go toolchain / gnark-crypto exists on the box, so the reference arm runs the C++ restatement
at the FULL size, one whole prove core per step, as many steps as fit in --reference-budget seconds.
"""
import argparse
import importlib.util
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "BN254 Groth16 prove s @2^24 constraints"
TOP_LIMB = 0x30644E72E131A029
R_MOD = 0x30644E72E131A029B85045B68181585D2833E84879B9709143E1F593F0000001
MAC32_PER_G1_POINT = 16 * 11 * 136   # 16 windows x (7M+4S) x 136 MAC32
MAC32_PER_G2_POINT = 16 * 29 * 136
MAC32_PER_G1_MADD = 8 * 136 + 2 * 64 + 72   # executed by XYZZ::add_mixed<Fp> (ec.cuh): 1288
MAC32_PER_BUTTERFLY = 136
# batched-affine round (msm_ba.cuh): 5 products + 1 square per addition, + the shared inversion: per 32 additions one thread
# product that costs 3 products in the inversion kernel and 1/128 of a 380-product Fermat chain
MAC32_PER_G1_AFFINE_ADD = 6 * 136 + (3 + 380 / 128.0) * 136 / 32.0


def g1_executed_mac32_per_point(info):
    """MAC32 the G1 bucket stage executes per point under the plan of one base set (b200_msm_set_info)."""
    left = 0.5 ** info.get("ba_rounds", 0)      # share of the additions left for the XYZZ kernel
    return info["windows"] * ((1.0 - left) * MAC32_PER_G1_AFFINE_ADD + left * MAC32_PER_G1_MADD)


def load_pkg():
    name = "gnark_fork_b200"
    if name in sys.modules:
        return sys.modules[name]
    d = os.path.join(ROOT, "gnark-fork_b200")
    spec = importlib.util.spec_from_file_location(name, os.path.join(d, "__init__.py"), submodule_search_locations=[d])
    m = importlib.util.module_from_spec(spec)
    sys.modules[name] = m
    spec.loader.exec_module(m)
    return m


def rand_field(rng, n, out=None, top_limb=TOP_LIMB):
    """n reduced field elements as Montgomery limbs (uniform up to the top-limb cut)."""
    a = out if out is not None else np.empty((n, 4), dtype=np.uint64)
    a[:, :3] = rng.integers(0, 1 << 64, size=(n, 3), dtype=np.uint64)
    a[:, 3] = rng.integers(0, top_limb, size=n, dtype=np.uint64)
    return a


def fr_limbs(x):
    return np.frombuffer(((x * (1 << 256)) % R_MOD).to_bytes(32, "little"), dtype=np.uint64).copy()


P_MOD = 0x30644E72E131A029B85045B68181585D97816A916871CA8D3C208C16D87CFD47
FR_ROOT_2_28 = 19103219067921713944291392827692070036145651957329286315305642004821462161904


def fp_limbs(v):
    return np.frombuffer(((v * (1 << 256)) % P_MOD).to_bytes(32, "little"), dtype=np.uint64).copy()


G1_GEN = np.concatenate([fp_limbs(v) for v in (1, 2)])
G2_GEN = np.concatenate([fp_limbs(v) for v in (
    10857046999023057135944570762232829481370756359578518086990519993285655852781,
    11559732032986387107991004021392285783925812861821192530917403151452391805634,
    8495653923123431417604973247489272438418190587263600148770280649306958101930,
    4082367875863433681332203403145435568316851327593401208105741076214120093531)])


def mont_sum(arr, r_mod=R_MOD):
    """Sum of Montgomery-form elements mod r, exact, vectorised (columns of 32-bit halves)."""
    h = np.ascontiguousarray(arr).view(np.uint32).reshape(-1, 8)
    tot = 0
    for j in range(8):
        tot += int(h[:, j].sum(dtype=np.uint64)) << (32 * j)
    return tot % r_mod


class CurveParams:
    """What the synthetic instance needs to know about a curve: Fr modulus and FFT root, generators as limbs."""

    def __init__(self, curve_id, name, r_mod, p_mod, fp_words, root, two_adicity, coset, g1, g2):
        self.id, self.name, self.r_mod, self.top_limb = curve_id, name, r_mod, r_mod >> 192
        self.root, self.two_adicity, self.coset = root, two_adicity, coset
        fp = lambda v: np.frombuffer(((v << (64 * fp_words)) % p_mod).to_bytes(8 * fp_words, "little"), dtype=np.uint64).copy()
        self.g1_gen = np.concatenate([fp(v) for v in g1])
        self.g2_gen = np.concatenate([fp(v) for v in g2])

    def fr_limbs(self, x):
        return np.frombuffer(((x << 256) % self.r_mod).to_bytes(32, "little"), dtype=np.uint64).copy()

    def generator(self, log_n):
        return self.fr_limbs(pow(self.root, 1 << (self.two_adicity - log_n), self.r_mod))


BN254_PARAMS = CurveParams(0, "bn254", R_MOD, P_MOD, 4, FR_ROOT_2_28, 28, 5, (1, 2), (
    10857046999023057135944570762232829481370756359578518086990519993285655852781,
    11559732032986387107991004021392285783925812861821192530917403151452391805634,
    8495653923123431417604973247489272438418190587263600148770280649306958101930,
    4082367875863433681332203403145435568316851327593401208105741076214120093531))
# BLS12-377 (config 5): r, p from std/math/emulated/emparams/emparams.go:123,137; Fr root of unity of order 2^47 =
# 22^((r-1)/2^47); coset shift 22; G1 generator as gnark-crypto's, G2 generator = a cofactor-cleared point (any
# generator does: the bases are inputs of the ABI) -- the same constants as oracle/bls12_377.py, which checks them.
_R377 = 0x12AB655E9A2CA55660B44D1E5C37B00159AA76FED00000010A11800000000001
_P377 = 0x01AE3A4617C510EAC63B05C06CA1493B1A22D9F300F5138F1EF3622FBA094800170B5D44300000008508C00000000001
BLS12_377_PARAMS = CurveParams(1, "bls12_377", _R377, _P377, 6, pow(22, (_R377 - 1) >> 47, _R377), 47, 22, (
    81937999373150964239938255573465948239988671502647976594219695644855304257327692006745978603320413799295628339695,
    241266749859715473739788878240585681733927191168601896383759122102112907357779751001206799952863815012735208165030), (
    67004216521362769139960304638241662463384504912045095543649118964167831559549275276665917802324363056069254155209,
    188039100033830534835560926333495938812099884491688240598849896699685545777288948680356069069856197422355451241144,
    212961125199539672632481864692443889005842964580434572770814582520396590909323644956646137861647360575214633605398,
    67625758907332218457243685014278362134590189520408250963023412301547932120327362287218456798771752606845621278268))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.proc, self.path, self.skip = device, None, None, 0

    def lines(self):
        try:
            return sum(1 for _ in open(self.path))
        except Exception:
            return 0

    def mark(self):
        """Samples taken so far (while the sampler spins up during warm-up) are dropped."""
        self.skip = self.lines()

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if not self.proc:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        try:
            for i, line in enumerate(open(self.path)):
                if i < self.skip:
                    continue
                p = [x.strip() for x in line.split(",")]
                if len(p) < 9:
                    continue
                try:
                    sm.append(float(p[1])); mx.append(float(p[2]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}
        return out


# ------------------------------------------------------------------------------------------
# the instance: refCircuit (groth16_test.go:120-156) with 2^log_n - 1 multiplications, its satisfying witness,
# key from b200_groth16_setup with a seeded toxic waste
# ------------------------------------------------------------------------------------------
class Instance:
    """nbConstraints = n = 2^log_n, wires = n + 2 (ONE, Y, X, X_1..X_{n-1}), A/B bases n each (2 infinity wires each),
    K = n private wires, Z = n - 1 (every MSM has ~2^log_n points).  The proving key comes out of the
    device Setup (this rank's point ranges); because tau, alpha, beta, gamma, delta are known, every MSM result and the
    proof have known discrete logs and the verify equation is checked in the exponent (groth16_host)."""

    def __init__(self, pkg, log_n, rank, world, seed=0xB200, cp=None):
        self.cp = cp = cp or BN254_PARAMS
        self.H = pkg.groth16_host
        self.fld = self.H.Field(cp.id)
        self.log_n, self.n = log_n, 1 << log_n
        n = self.n
        self.n_wires, self.nb_public = n + 2, 2
        self.full = dict(a=n, b=n, k=n, z=n - 1)
        self.rank, self.world = rank, world
        self.gen = cp.generator(log_n)
        self.coset = cp.fr_limbs(cp.coset)
        self.toxic = self.H.sample_toxic(self.fld, np.random.default_rng(seed + log_n))   # same on every rank

    def build_key(self, ctx):
        cs, pk, res = self.H.setup_mul_chain(ctx, self.fld, self.log_n, self.toxic, self.cp.g1_gen, self.cp.g2_gen,
                                             shard_index=self.rank, shard_count=self.world)
        assert (res["n_a"], res["n_b"], res["n_k"], res["n_z"]) == tuple(self.full[k] for k in "abkz")
        self.vk = res
        return pk

    def witness(self, ctx, pkg):
        """W, a, b, c of the chain X <- X * X from X = 2 in page-locked host memory."""
        self.w, self.a, self.b, self.c = self.H.mul_chain_witness(ctx, self.fld, self.log_n, 2, alloc=pkg.pinned_empty)

    def vk_points(self):
        return [self.vk[k] for k in ("g1_alpha", "g1_beta", "g1_delta", "g2_beta", "g2_delta")]

    def check(self, ctx, pkg, total, raw_out=None):
        """`total` = the five MSM results summed over the ranks (dict of Jacobian points); raw_out = the b200_groth16_out
        they came from when there is one rank (then the host fix-ups are checked too).  Exact at any size: each must be [e]G for the discrete
        log e the satisfying witness implies (A(tau), B(tau), K and (A B - C)(tau) / delta -- the device's h does not enter),
        and the proof built from them must satisfy verify.go:65-137 in the exponent."""
        H = self.H
        exp = H.mul_chain_expected_dlogs(ctx, self.fld, self.log_n, self.toxic, self.w, self.a, self.b, self.c)
        ok = H.points_match(ctx, pkg, self.fld, self.cp.g1_gen, self.cp.g2_gen, total, exp)
        if not all(ok.values()):
            raise SystemExit(f"bench: MSM results do not match the witness's discrete logs {ok} -- result invalid")
        r_, s_ = 0x1234567 % self.fld.r, 0x7654321 % self.fld.r
        proof = H.expected_proof_dlogs(self.fld, self.toxic, exp, r_, s_)
        if not H.verify_equation(self.fld, self.toxic, exp, proof):
            raise SystemExit("bench: the verify equation does not hold -- result invalid")
        if raw_out is not None:     # the finalised proof points are the expected ones too (prove.go:178-321 on the host)
            ar, bs, krs = pkg.groth16_finalize(raw_out, self.fld.limbs(r_), self.fld.limbs(s_), *self.vk_points(), curve=self.cp.id)
            want_ar = ctx.batch_scalar_mul(self.cp.g1_gen, self.fld.limbs(proof["ar"]).reshape(1, 4), 1)[0]
            want_krs = ctx.batch_scalar_mul(self.cp.g1_gen, self.fld.limbs(proof["krs"]).reshape(1, 4), 1)[0]
            want_bs = ctx.batch_scalar_mul(self.cp.g2_gen, self.fld.limbs(proof["bs"]).reshape(1, 4), 2)[0]
            if not (np.array_equal(ar, want_ar) and np.array_equal(krs, want_krs) and np.array_equal(bs, want_bs)):
                raise SystemExit("bench: finalised proof differs from the expected group elements -- result invalid")
        return True


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle restatement (oracle/cpu_ref.cpp) on all host cores.  No Go toolchain / gnark-crypto exists on
# the box, so this C++ restatement stands for the reference's CPU prover ("kind": "port").
# ------------------------------------------------------------------------------------------
def workload_config(log_n, world):
    """The `config` object of BOTH arms (the reference arm runs on this arm's config)."""
    return {"workload": f"BN254 Groth16 prove core (computeH + 5 MSM) of the refCircuit Mul chain (groth16_test.go:120-156), "
                        f"2^{log_n} constraints, satisfying witness (4 G1 MSM of ~2^{log_n}, 1 G2 MSM, 7 NTT of 2^{log_n})",
            "log_n": log_n, "inputs": "X <- X*X chain from X = 2; proving key from a seeded toxic waste",
            "parallelism": (f"msm point-range shards x{world}, computeH distributed (one all-to-all per transform)" if world > 1 else "1 GPU"),
            "l2": "inputs (>= 0.5 GiB per vector) exceed the 126 MB L2; no flush needed"}


def cpu_threads():
    from oracle import cref
    try:
        avail = len(os.sched_getaffinity(0))      # torchrun exports OMP_NUM_THREADS=1: not what is wanted here
    except AttributeError:
        avail = os.cpu_count() or 1
    cref.BN254.set_threads(int(avail))
    return cref.BN254.max_threads()


def cpu_chain_witness(log_n):
    """(w, a, b, c) of the chain in Montgomery form, made on the CPU (oracle fr_mul by R^2 for the conversion)."""
    from oracle import cref
    m = 1 << log_n
    nb_mul = m - 1
    buf = bytearray(32 * m)
    cur = 2
    buf[0:32] = cur.to_bytes(32, "little")
    for j in range(1, m):
        cur = cur * cur % R_MOD
        buf[32 * j:32 * j + 32] = cur.to_bytes(32, "little")
    canon = np.frombuffer(bytes(buf), dtype=np.uint64).reshape(-1, 4)
    r2 = np.frombuffer(((1 << 512) % R_MOD).to_bytes(32, "little"), dtype=np.uint64)
    chain = cref.fr_mul(canon, np.broadcast_to(r2, canon.shape).copy())
    one = fr_limbs(1)
    w = np.concatenate([one[None], chain[nb_mul:nb_mul + 1], chain])
    a = np.concatenate([chain[:nb_mul], one[None]])
    b = np.concatenate([chain[:nb_mul], chain[nb_mul:]])
    c = np.concatenate([chain[1:], chain[nb_mul:]])
    return w, a, b, c


def cpu_prove_once(log_n, w, a, b, c, g1_sets, g2_b):
    """One prove core of the C++ restatement at 2^log_n: computeH (7 transforms) + MultiExp over pk.G1.A, B, K, Z and
    pk.G2.B with the scalars prove.go:201-290 pairs them with.  Returns (seconds, {part: seconds}, results)."""
    from oracle import cref
    n = 1 << log_n
    gen = fr_limbs(pow(FR_ROOT_2_28, 1 << (28 - log_n), R_MOD))
    t0 = time.perf_counter()
    h = cref.compute_h(a, b, c, log_n, gen, fr_limbs(5))
    t_h = time.perf_counter() - t0
    wa = np.ascontiguousarray(np.concatenate([w[:1], w[2:n + 1]]))     # InfinityA: Y and X_N
    wb = np.ascontiguousarray(w[2:])                                    # InfinityB: ONE and Y
    wk = wb                                                             # W[nb_public:]
    t1 = time.perf_counter()
    res = dict(ar=cref.msm_g1(wa, g1_sets["a"]), bs1=cref.msm_g1(wb, g1_sets["b"]), krs=cref.msm_g1(wk, g1_sets["k"]),
               krs2=cref.msm_g1(np.ascontiguousarray(h[:n - 1]), g1_sets["z"]))
    t_g1 = time.perf_counter() - t1
    t2 = time.perf_counter()
    res["bs2"] = cref.msm_g2(wb, g2_b)
    t_g2 = time.perf_counter() - t2
    return time.perf_counter() - t0, dict(compute_h=t_h, msm_g1_x4=t_g1, msm_g2=t_g2), res, h


def cpu_synthetic_key(log_n, distinct_log=18):
    """Base arrays for the CPU-only arm.  The cost of Pippenger does not depend on which points the bases are, so the
    arrays are 2^distinct_log distinct multiples of the generator (oracle fixed-base multiplication) tiled to length."""
    from oracle import cref
    n = 1 << log_n
    m = 1 << min(distinct_log, log_n)
    k = rand_field(np.random.default_rng(7), m)
    g1 = cref.g1_batch_mul(G1_GEN, k)
    g2 = cref.g2_batch_mul(G2_GEN, k)
    reps = n // m
    t1 = np.tile(g1, (reps, 1))
    return dict(a=t1, b=t1, k=t1, z=t1[:n - 1]), np.tile(g2, (reps, 1))


def cpu_baseline_leg(ctx, pkg, log_n, sample_log):
    """cpu_baseline of the GPU arm's line: ONE prove core of the C++ restatement at 2^sample_log on the same kind of
    instance (device Setup with the points copied back, the chain's witness), timed on all host cores, and -- the
    checker doing its job -- its five MSM results and its h compared with the device's on the same inputs.  The
    figure is scaled to 2^log_n: MSM part linearly (the restatement's window is 16 bits from 2^19 points up, as
    gnark-crypto's largest), transform part by n log n."""
    cores = cpu_threads()
    H = pkg.groth16_host
    fld = H.Field(pkg.BN254)
    toxic = H.sample_toxic(fld, np.random.default_rng(99))
    cs, pk, res = H.setup_mul_chain(ctx, fld, sample_log, toxic, G1_GEN, G2_GEN, want_points=True)
    w, a, b, c = H.mul_chain_witness(ctx, fld, sample_log, 2)
    dev_out, _ = ctx.groth16_prove_core(pk, a, b, c, w)
    dev_h = ctx.compute_h(a, b, c, sample_log, fld.limbs(fld.generator(sample_log)), fld.limbs(fld.coset))
    pk.free()
    sec, parts, cpu_res, cpu_h = cpu_prove_once(sample_log, w, a, b, c, dict(a=res["g1_a"], b=res["g1_b"], k=res["g1_k"], z=res["g1_z"]), res["g2_b"])
    from oracle import cref
    od = dev_out.as_dict()
    same = bool(np.array_equal(cpu_h, dev_h))
    for k in ("ar", "bs1", "krs", "krs2"):
        same = same and bool(np.array_equal(cref.g1_jac_to_aff(cpu_res[k]), pkg.g1_jac_to_affine(od[k])))
    same = same and bool(np.array_equal(cref.g2_jac_to_aff(cpu_res["bs2"]), pkg.g2_jac_to_affine(od["bs2"])))
    if not same:
        raise SystemExit("bench: the CPU restatement and the device disagree on the sample instance -- result invalid")
    up = 1 << (log_n - sample_log)
    est = (parts["msm_g1_x4"] + parts["msm_g2"]) * up + parts["compute_h"] * up * log_n / sample_log
    text = (f"one prove core (computeH + 4 G1 MSM + 1 G2 MSM) of the C++ restatement on the refCircuit instance at 2^{sample_log} "
            f"constraints: {sec:.2f} s on {cores} threads (computeH {parts['compute_h']:.2f}, G1 x4 {parts['msm_g1_x4']:.2f}, G2 {parts['msm_g2']:.2f}); "
            f"h and all five MSM results equal the device's on the same inputs; scaled to 2^{log_n}: MSM x{up} (fixed 16-bit window), "
            f"transforms x{up} x {log_n}/{sample_log}.  The full-size run is the reference arm (--impl reference).  "
            "CPU restatement, not gnark-crypto (no Go toolchain on the box, no amd64 assembly)")
    return {"value": est, "unit": "s", "cores": cores, "kind": "port", "sample": text, "sample_s": sec, "sample_log_n": sample_log,
            "sample_matches_device": same}


def reference_arm(args):
    """--impl reference: the reference's CPU path for this metric, measured -- not extrapolated -- at the full size.
    The reference itself (Go, gnark-crypto) cannot run here; its C++ restatement runs ONE whole prove core per step at
    2^log_n constraints on all host cores.  A step takes about a minute, so the arm runs as many of the requested
    steps as fit in --reference-budget seconds (at least one) and reports the count it ran; the warm-up is one small
    untimed run (thread pool, page faults), reported as warmup 0 of the measured size."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = cpu_threads()
    log_n = args.log_n
    t_setup = time.perf_counter()
    g1_sets, g2_b = cpu_synthetic_key(log_n)
    w, a, b, c = cpu_chain_witness(log_n)
    small = cpu_chain_witness(12)
    ks, k2 = cpu_synthetic_key(12)
    cpu_prove_once(12, *small, ks, k2)
    t_setup = time.perf_counter() - t_setup
    times, parts_all = [], []
    t_begin = time.perf_counter()
    while len(times) < max(1, args.steps):
        sec, parts, _, _ = cpu_prove_once(log_n, w, a, b, c, g1_sets, g2_b)
        times.append(sec)
        parts_all.append(parts)
        if time.perf_counter() - t_begin + sec > args.reference_budget:
            break
    v = float(np.mean(times))
    parts = {k: float(np.mean([p[k] for p in parts_all])) for k in parts_all[0]}
    text = (f"{len(times)} whole prove core(s) of the C++ restatement at 2^{log_n} constraints, measured (no extrapolation): "
            f"computeH {parts['compute_h']:.1f} s, 4 G1 MSM {parts['msm_g1_x4']:.1f} s, G2 MSM {parts['msm_g2']:.1f} s on {cores} threads; "
            f"witness = the refCircuit chain, bases = 2^18 distinct points tiled (Pippenger's cost does not depend on the points); "
            f"of the {args.steps} steps requested, as many as fit in {args.reference_budget:.0f} s were run.  "
            "CPU restatement, not gnark-crypto (no Go toolchain on the box, no amd64 assembly)")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "s", "n_gpus": args.gpus, "steps": len(times),
            "warmup": 0, "ms_per_step": v * 1e3, "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": "u32 limbs (254-bit Montgomery)", "data": "synthetic",
            "config": workload_config(log_n, args.gpus),
            "cpu_baseline": {"value": v, "unit": "s", "cores": cores, "kind": "port", "sample": text, "stage_s": parts,
                             "steps_requested": args.steps, "setup_s": t_setup},
            "e2e": {"value": v, "unit": "s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
def _max_over_ranks(dist, local, x):
    import torch
    t = torch.tensor([float(x)], dtype=torch.float64, device=f"cuda:{local}")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gpu_arm(args):
    """N = 1: one context.  N > 1: ONE process proves on all N GPUs through the device group (b200_group_*: host thread
    per rank inside the library, sharded MSMs combined in C, computeH's exchanges fused into the kernels over NVLink peer
    memory).  Under torchrun (how the driver launches N > 1) rank 0 is that process; the other ranks join the NCCL
    communicator, then wait on a CPU (gloo) barrier so that no spinning collective kernel sits on the GPUs rank 0 is
    timing.  The PlonK (config 4) and BLS12-377 (config 5) sections run on the same N GPUs through their own groups."""
    rank = int(os.environ.get("RANK", "0"))
    launched = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = launched if launched > 1 else max(1, args.gpus)       # plain `python bench.py --gpus N` also works: one process
    dist = gloo = None
    if launched > 1:
        import datetime
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        # NCCL prints its version banner on stdout while the communicator comes up: send fd 1 to
        # stderr for that moment so that stdout carries the one JSON line only
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(hours=2))
            dist.barrier()
            torch.cuda.synchronize()
            gloo = dist.new_group(backend="gloo", timeout=datetime.timedelta(hours=2))
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    pkg = load_pkg()
    if launched > 1 and rank != 0:
        dist.barrier(group=gloo)          # rank 0 has finished: every section runs from its one process
        dist.destroy_process_group()
        return 0

    grp = None
    if world == 1:
        ctx = pkg.Context(local)          # raises without the CUDA library / device: no fallback
    else:
        grp = pkg.Group([int(x) for x in args.devices.split(",")] if args.devices else list(range(world)))
        ctx = grp.ctx(0)
    peak = ctx.imad_peak()
    for kv in args.opt:
        k, v = kv.split("=")
        (grp or ctx).set_option(k, int(v))
    inst = Instance(pkg, args.log_n, 0, 1)
    t_setup = time.perf_counter()
    pk = inst.build_key(grp or ctx)
    t_key = time.perf_counter() - t_setup
    inst.witness(ctx, pkg)
    t_setup = time.perf_counter() - t_setup
    n, nw = inst.n, inst.n_wires
    rs = rand_field(np.random.default_rng(3), 2)
    vkp = inst.vk_points()

    if grp is None:
        distributed = False
        dv = {k: pkg.DeviceVector(ctx, getattr(inst, k).nbytes).upload(getattr(inst, k)) for k in ("a", "b", "c", "w")}

        def step_dev():
            o, tm = ctx.groth16_prove_core_dev(pk, dv["a"], dv["b"], dv["c"], n, dv["w"])
            return o, tm, tm["total_ms"]

        def step_e2e():
            return ctx.groth16_prove_core(pk, inst.a, inst.b, inst.c, inst.w)[0]
        w_lo, w_hi = pk.wire_range()
        h2d_bytes = int((3 * n + (w_hi - w_lo)) * 32)
        pk_info = pk.info()
    else:
        pk_info = pk.info(0)
        distributed = pk_info.pop("distributed_h")
        ctxs = [grp.ctx(r) for r in range(world)]
        h2d_bytes = int((3 * n + nw + 2 * (world - 1)) * 32)      # a, b, c blocks + each rank's wire range of W (ranges overlap by a wire or two)
        if distributed:
            n_loc = n // world
            blocks = [[pkg.DeviceVector(ctxs[r], n_loc * 32).upload(np.ascontiguousarray(getattr(inst, k)[r * n_loc:(r + 1) * n_loc]))
                       for r in range(world)] for k in ("a", "b", "c")]
            wv = [pkg.DeviceVector(ctxs[r], inst.w.nbytes).upload(inst.w) for r in range(world)]

            def step_dev():
                o, tm = grp.groth16_prove_core_dev(pk, blocks[0], blocks[1], blocks[2], wv)
                return o, tm, tm["total_ms"]
        else:       # odd group size: every rank runs the whole computeH from the host buffers
            def step_dev():
                o, tm = grp.groth16_prove_core(pk, inst.a, inst.b, inst.c, inst.w)
                return o, tm, tm["total_ms"]

        def step_e2e():
            return grp.groth16_prove_core(pk, inst.a, inst.b, inst.c, inst.w)[0]

    def barrier():
        for c in ([ctx] if grp is None else ctxs):
            c.sync()

    # correctness first (outside any timed region): the proof of the satisfying witness must verify -- all five MSM
    # results against their discrete logs, the verify equation in the exponent, the host fix-ups (Instance.check)
    out, _, _ = step_dev()
    verified = None if args.no_check else inst.check(ctx, pkg, out.as_dict(), out)
    if not args.no_check and grp is not None:       # and the host-buffer path of the group
        inst.check(ctx, pkg, step_e2e().as_dict())

    clocks = ClockSampler(local)
    clocks.start()                      # nvidia-smi needs a moment to come up: start it under the warm-up
    for _ in range(max(0, args.warmup - 1)):
        step_dev()

    # ---- device-resident timed region: K steps, CUDA events (max over the ranks, taken inside the library) ----
    barrier()
    clocks.mark()
    ctx.profile_start()
    dev_ms, stage = [], {}
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _, tm, total = step_dev()
        dev_ms.append(total)
        for k, v in tm.items():
            stage[k] = stage.get(k, 0.0) + v / args.steps
    prof = ctx.profile_read()
    barrier()
    wall_dev = time.perf_counter() - t0
    extra = 0
    while extra < 200 and clocks.lines() - clocks.skip < 3 and time.perf_counter() - t0 < 5.0:
        step_dev()                      # region shorter than the 100 ms sampling period: same load, untimed
        extra += 1
    clk = clocks.stop()
    clk["window"] = "timed region" if not extra else f"timed region + {extra} identical untimed steps (region shorter than the sampling period)"
    step_ms = float(np.mean(dev_ms))
    stage["total_ms"] = step_ms

    # ---- end to end through the host-buffer path: pinned host -> device(s), MSM results back (combined),
    #      host fix-ups (prove.go:178-321) to the affine proof -------------------------------
    step_e2e()
    barrier()
    t0 = time.perf_counter()
    fin_s = 0.0
    for _ in range(args.steps):
        o = step_e2e()
        t1 = time.perf_counter()
        pkg.groth16_finalize(o, rs[0], rs[1], *vkp)
        fin_s += time.perf_counter() - t1
    barrier()
    e2e_s = (time.perf_counter() - t0) / args.steps
    args.finalize_ms = fin_s / args.steps * 1e3

    args.plonk_result = None
    pk.free()                           # the folded tables go before the PlonK / BLS12-377 keys come
    if grp is not None:
        for v in [x for vs in (blocks if distributed else []) for x in vs] + (wv if distributed else []):
            v.free()
        grp.close()
        ctx = pkg.Context(local)
    args.bls_result = None
    if args.bls12_377_log_n:            # config 5, on the same number of GPUs (its own library and device group)
        try:
            devs = ([int(x) for x in args.devices.split(",")] if args.devices else list(range(world))) if world > 1 else None
            args.bls_result = bls12_377_section(pkg, local, args.bls12_377_log_n, args.steps, args.warmup, peak, devs)
        except Exception as e:          # the headline line must still be printed
            args.bls_result = {"error": repr(e)}
    args.cpu_result = None
    if world == 1 and not args.no_cpu:  # the one leg that runs the oracle: the CPU baseline (and the device against it)
        args.cpu_result = cpu_baseline_leg(ctx, pkg, inst.log_n, min(args.cpu_sample_log, inst.log_n))
    if args.plonk_log_n and (world == 1 or args.plonk_dist):     # config 4 on the same number of GPUs
        try:
            devs = ([int(x) for x in args.devices.split(",")] if args.devices else list(range(world))) if world > 1 else None
            args.plonk_result = plonk_section(pkg, ctx, args.plonk_log_n, args.steps, args.warmup, devs)
        except Exception as e:          # the headline line must still be printed
            args.plonk_result = {"error": repr(e)}
    if launched > 1:
        dist.barrier(group=gloo)
        dist.destroy_process_group()
    return report(args, world, peak, prof, stage, step_ms, e2e_s, clk, verified, inst, t_setup, wall_dev, pk_info, distributed, h2d_bytes)


def bls12_377_section(pkg, device_index, log_n, steps, warmup, peak, devices=None):
    """BASELINE.json config 5: BLS12-377 Groth16 prove core at 2^log_n constraints on one GPU through the second
    library (libb200prover_bls12377.so: 12-limb Fp, Fp2 with u^2 = -5; same kernels, same ABI).  Same instance shape
    and the same exact check as the BN254 headline (every MSM result against its known discrete log).  `value` =
    seconds per proof through the host-buffer call (pinned a, b, c, W in; H2D inside) + the host fix-ups."""
    grp = pkg.Group(devices, curve=pkg.BLS12_377) if devices and len(devices) > 1 else None     # N > 1: the device group of the second library
    ctx = grp.ctx(0) if grp else pkg.Context(device_index, curve=pkg.BLS12_377)
    prover = grp or ctx
    world = len(devices) if grp else 1
    try:
        inst = Instance(pkg, log_n, 0, 1, cp=BLS12_377_PARAMS)
        t0 = time.perf_counter()
        pk = inst.build_key(prover)
        inst.witness(ctx, pkg)
        t_setup = time.perf_counter() - t0
        rs = rand_field(np.random.default_rng(3), 2, top_limb=inst.cp.top_limb)
        vkp = inst.vk_points()
        for _ in range(max(1, warmup)):
            out, tm = prover.groth16_prove_core(pk, inst.a, inst.b, inst.c, inst.w)
        ctx.profile_start()
        wall, tms = [], []
        for _ in range(steps):
            t1 = time.perf_counter()
            out, tm = prover.groth16_prove_core(pk, inst.a, inst.b, inst.c, inst.w)
            pkg.groth16_finalize(out, rs[0], rs[1], *vkp, curve=pkg.BLS12_377)
            wall.append(time.perf_counter() - t1)
            tms.append(tm)
        prof = ctx.profile_read()
        ok = inst.check(ctx, pkg, out.as_dict(), out)
        info = pk.info(0) if grp else pk.info()
        pk.free()
        n = inst.n
        g1 = prof["msm_acc_g1"]
        g1_launch_ms = g1["ms"] / max(1, g1["launches"])
        mac_fp = 12 * 12 * 2 + 12                 # one 12-limb Montgomery product: 288 + 12 MAC32 (300)
        pts = g1["units"] / max(1, g1["launches"])
        model = pts * 16 * 11 * mac_fp / (g1_launch_ms * 1e-3) if g1_launch_ms else 0.0
        mac_madd = 8 * mac_fp + 2 * 144 + 156     # XYZZ mixed addition: 8 products + a*b - c*d as two wide products and ONE reduction
        ba_r = info["a"].get("ba_rounds", 0)
        left = 0.5 ** ba_r
        mac_aff = 6 * mac_fp + (3 + 570 / 128.0) * mac_fp / 32.0      # affine addition with shared inversions (Fermat chain of a 377-bit p: ~570 products)
        executed = pts * info["a"]["windows"] * ((1 - left) * mac_aff + left * mac_madd) / (g1_launch_ms * 1e-3) if g1_launch_ms else 0.0
        stage = {k: float(np.mean([t[k] for t in tms])) for k in tms[0]}
        return {"metric": "BLS12-377 Groth16 prove s @2^%d constraints" % log_n, "value": float(np.mean(wall)), "unit": "s", "n_gpus": world,
                "steps": steps, "workload": "synthetic 2^%d-constraint Mul-chain-shaped R1CS on BLS12-377: computeH (7 transforms of 2^%d) + "
                "4 G1 MSMs + 1 G2 MSM of ~2^%d points (Fp 377 bits = 12 x 32-bit limbs)" % (log_n, log_n, log_n),
                "path": ("b200_groth16_prove_core" if not grp else "b200_group_groth16_prove_core over %d GPUs" % world) +
                        " of libb200prover_bls12377.so (pinned host a, b, c, W in) + b200_groth16_finalize",
                "device_ms": stage["total_ms"], "stage_ms": stage, "verified": bool(ok),
                "verifier": "refCircuit with its satisfying witness, key from b200_groth16_setup (known toxic waste): every MSM result equals "
                            "[e]G for the discrete log e the witness implies, the verify equation holds in the exponent, host fix-ups checked",
                "roofline_g1": {"bound": "imad", "kernel": "G1 bucket stage, 12 limbs: %d batched-affine rounds + msm_accumulate_kernel<Fp>" % ba_r, "launch_ms": g1_launch_ms,
                                "model_mac32_per_s": model, "executed_mac32_per_s": executed, "peak_mac32_per_s": peak["mac32_per_s"],
                                "frac_model": model / peak["mac32_per_s"], "frac_executed": executed / peak["mac32_per_s"],
                                "note": "model = points x 16 windows x 11 Fp products x 300 MAC32 (SURVEY.md 8(d)); executed = points x the plan's "
                                        "windows x [(1 - 2^-rounds) affine additions with shared inversions (6 products x 300 + the inversion share) + 2^-rounds XYZZ mixed "
                                        "additions (2844 MAC32: 8 products x 300 + two 144-MAC32 wide products + one 156-MAC32 reduction)]; launch_ms is the stage's span "
                                        "inside the prove core, beside the next MSM's sort"},
                "msm_plan": {k: info["a"][k] for k in ("c", "windows", "folded", "ba_rounds")},
                "gpu_launches_per_proof": int(prof["kernel_launches"] // max(1, steps)), "setup_s": t_setup}
    finally:
        (grp or ctx).close()


def plonk_section(pkg, ctx, log_n, steps, warmup, devices=None):
    """BASELINE.json config 4: BN254 PlonK Prove at 2^log_n gates (SCS Mul chain, backend/plonk/plonk_test.go:248-260
    shape) with page-locked host L, R, O; every timed proof is checked with the trapdoor form of plonk.Verify
    (verify.go:45-294; plonk_host.verify_with_trapdoor).  Wall time per proof, host buffers in, proof out.
    N = 1: b200_plonk_prove.  N > 1: b200_group_plonk_prove from this one process -- the key sharded over the group
    (commitments by SRS point range, quotient by coset), the exchanges inside the library."""
    import importlib
    H = importlib.import_module("gnark_fork_b200.plonk_host")
    n = 1 << log_n
    world = len(devices) if devices else 1
    tau = 0x1234567890ABCDEF1234567890ABCDEF % H.R_MOD
    t0 = time.perf_counter()
    trace, lro, nb_vars, nb_pub, L, Rr, O, pub = H.mul_chain(log_n, 3)
    d = H.Domains(n)
    kw = H.setup(ctx, d, trace, lro, nb_vars, nb_pub, tau)
    grp = pkg.Group(devices) if world > 1 else None
    prover = grp or ctx
    dpk = prover.plonk_pk_upload(**kw)
    t_setup = time.perf_counter() - t0
    bufs = []
    for a in (L, Rr, O):
        b = pkg.pinned_empty(a.shape)
        b[:] = a
        bufs.append(b)
    c0 = grp.ctx(0) if grp else ctx

    def one():
        return prover.plonk_prove(dpk, *bufs, pub, None)

    c0.profile_start()
    for _ in range(max(1, warmup)):
        one()
    launches = c0.profile_read()["kernel_launches"] // max(1, warmup)
    wall, stages = [], []
    for _ in range(steps):
        t1 = time.perf_counter()
        proof, tm = one()
        wall.append((time.perf_counter() - t1) * 1e3)
        stages.append(tm)
    ms = float(np.mean(wall))
    ok = H.verify_with_trapdoor(proof, kw["vk_commitments"], H.from_mont_array(pub), d, tau)
    dpk.free()
    if grp:
        grp.close()
    mean_stage = {k: float(np.mean([s[k] for s in stages])) for k in stages[0]}
    msm_ms = sum(mean_stage[k] for k in ("commit_lro_ms", "commit_z_ms", "commit_h_ms", "batch_open_ms")) + mean_stage["linearize_ms"]
    return {"metric": "BN254 PlonK prove ms @2^%d gates" % log_n, "value": ms, "unit": "ms", "n_gpus": world, "steps": steps,
            "workload": "SCS Mul chain filling 2^%d rows, big domain 2^%d; 9 KZG commitments as G1 MSMs of ~2^%d against 2 resident SRS (+ the linearised digest as a 7-point host combination), "
                        "%d transforms of 2^%d + 1 of 2^%d, fresh blinding per proof" % (log_n, d.log_big, log_n, 5 + 5 * (d.big // n), log_n, d.log_big),
            "path": "b200_plonk_prove (page-locked host L, R, O in; proof out; Fiat-Shamir on the host)" if world == 1 else
                    "b200_group_plonk_prove, one process, %d ranks: commitments sharded by SRS point range (partial sums added on the host inside the call), "
                    "quotient sharded by coset (numerator blocks by peer copies)" % world,
            "h2d_bytes_per_step": int(3 * n * 32 + 32) * world, "d2h_bytes_per_step": 552,
            "stage_ms": mean_stage, "stage_ms_note": "rank 0, library time only",
            "commit_ms_share": msm_ms / mean_stage["total_ms"],
            "gpu_launches_per_proof": int(launches), "verified": bool(ok), "verifier": "plonk.Verify with the KZG pairing checks done in G1 through the known tau of the test SRS",
            "setup_s": t_setup}


def report(args, world, peak, prof, stage, step_ms, e2e_s, clk, verified, inst, t_setup, wall_dev, pk_info, distributed, h2d_bytes):
    n = inst.n
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        hbm_peak, hbm_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    else:
        hbm_peak, hbm_src = 6650.0, "fallback (B200_PROFILING.md)"
    mac_peak = peak["mac32_per_s"]
    g1, g2, ntt, pw = prof["msm_acc_g1"], prof["msm_acc_g2"], prof["ntt_pass"], prof["pointwise"]
    # dominant kernel: msm_accumulate_kernel<Fp>; algorithmic MAC32 per launch = points x 23 936
    g1_launch_ms = g1["ms"] / max(1, g1["launches"])
    g1_pts = g1["units"] / max(1, g1["launches"])
    g1_mac_s = g1_pts * MAC32_PER_G1_POINT / (g1_launch_ms * 1e-3)
    ba_rounds = pk_info["a"].get("ba_rounds", 0)
    exec_pp = g1_executed_mac32_per_point(pk_info["a"])
    if ba_rounds:
        kernel = ("G1 bucket stage = %d batched-affine rounds (ba_apply_kernel<Fp> + ba_prep_kernel<Fp> + ba_inverse_kernel<Fp>; the first "
                  "round's ba_prep_kernel runs with the sort) + msm_accumulate_kernel<Fp> on the 1/%d of the additions that are left" % (ba_rounds, 1 << ba_rounds))
        # ncu dram__bytes_read.sum + dram__bytes_write.sum over the stage's launches at 2^24 points, 4 rounds (profiles/r02_ba_launches_g1_2p24.csv)
        traffic = 78.0e9 * g1_pts / 16777216.0 if (ba_rounds == 4 and pk_info["a"]["c"] == 20) else None
        traffic_source = ("ncu dram__bytes_read.sum + dram__bytes_write.sum summed over the stage's 13 launches at 2^24 points, c = 20, 4 rounds "
                          "(profiles/r02_ba_launches_g1_2p24.csv: 57.6 GB read + 20.4 GB written), scaled to this launch's points; the table "
                          "gathers fetch 64-byte halves (ld.global.nc.L2::64B, profiles/r02_gather_probe.txt)")
        exec_model = (f"{pk_info['a']['windows']} windows of c={pk_info['a']['c']} x [(1 - 1/{1 << ba_rounds}) affine additions with shared inversions "
                      f"(5 products + 1 square + 0.19 of the inversion kernel = {MAC32_PER_G1_AFFINE_ADD:.0f} MAC32) + 1/{1 << ba_rounds} XYZZ mixed additions (1288 MAC32)]; "
                      "frac > 1 means fewer field multiplications than the survey's model, not a faster pipe")
    else:
        kernel = "msm_accumulate_kernel<Fp> (G1 bucket accumulation)"
        traffic = (3.77e9 * g1_pts / 4194304.0) if pk_info["a"]["folded"] and pk_info["a"]["c"] == 20 else None
        traffic_source = ("ncu --set full at 2^22 points (profiles/r01_prof_fold_2p22_raw.csv: 7.54 GB with 128-byte line fetches), halved by the "
                          "64-byte gather hint of round 2 (profiles/r02_gather_probe.txt), scaled linearly to this launch's points")
        exec_model = (f"{pk_info['a']['windows']} windows of c={pk_info['a']['c']} "
                      f"({'folded table' if pk_info['a']['folded'] else 'plain bases'}) x XYZZ mixed add: 8 Montgomery products x 136 "
                      "+ one a*b - c*d with a shared reduction (2 x 64 + 72) = 1288 MAC32; "
                      "frac > 1 means fewer field multiplications than the survey's model, not a faster pipe")
    roofline = {"kernel": kernel, "bound": "imad",
                "achieved": g1_mac_s / 1e12, "peak": mac_peak / 1e12, "unit": "TMAC32/s", "frac": g1_mac_s / mac_peak,
                "traffic": traffic, "traffic_source": traffic_source,
                "hbm_frac_of_traffic": (traffic / (g1_launch_ms * 1e-3) / 1e9 / hbm_peak) if traffic and g1_launch_ms else None,
                "launch_ms": g1_launch_ms, "points_per_launch": g1_pts,
                "launch_ms_note": "CUDA-event span of the stage on the main stream inside the timed region; another MSM's sort runs beside it on the aux stream and stretches it (alone: 24.1 ms at 2^24 points)",
                "algorithmic_mac32_per_point": MAC32_PER_G1_POINT,
                "algorithmic_model": "SURVEY 8(d): 16 windows x 11 Fp-mul x 136 MAC32 per point, whatever the kernel executes",
                "executed_mac32_per_point": exec_pp,
                "executed_frac": g1_pts * exec_pp / (g1_launch_ms * 1e-3) / mac_peak,
                "executed_model": exec_model,
                "peak_source": "measured in this run: register-resident Montgomery-product chains (b200_imad_peak)"}
    # NTT: north_star's HBM view (64 B per element per pass, 2 passes at <= 2^24) and the IMAD view
    passes_model = 1 if inst.log_n <= 12 else (2 if inst.log_n <= 24 else 3)
    n_transforms = 7 * args.steps
    ntt_bytes = n_transforms * (n / world if distributed else n) * 64.0 * passes_model   # per rank
    ntt_gbs = ntt_bytes / (ntt["ms"] * 1e-3) / 1e9 if ntt["ms"] else 0.0
    ntt_mac_s = ntt["units"] * MAC32_PER_BUTTERFLY / (ntt["ms"] * 1e-3) if ntt["ms"] else 0.0
    # measured DRAM traffic of one 2^24 transform (ncu --set full, profiles/r02_final_ntt_2p24_ncu.txt): forward DIF 2.55 GB read +
    # 1.46 GB written, inverse DIT on the coset (scale table on the last store) 3.11 + 1.47 GB, in 3 pass launches each
    ntt_traffic_per_transform = 0.5 * (4.01e9 + 4.58e9) * ((n / world if distributed else n) / float(1 << 24))
    roofline_ntt = {"kernel": "ntt_pass_tma_kernel", "bound": "hbm", "achieved": ntt_gbs, "peak": hbm_peak, "unit": "GB/s",
                    "frac": ntt_gbs / hbm_peak, "traffic": ntt_traffic_per_transform,
                    "traffic_source": "ncu dram__bytes_read.sum + dram__bytes_write.sum of the 3 pass launches of one 2^24 transform (mean of a forward DIF, 4.01 GB, "
                                      "and an inverse coset DIT, 4.58 GB; profiles/r02_final_ntt_2p24_ncu.txt), per transform, scaled linearly to this size; "
                                      "algorithmic: 64 B per element per pass = %.2f GB per transform at 2 passes" % ((n / world if distributed else n) * 64.0 * passes_model / 1e9),
                    "peak_source": hbm_src,
                    "imad_frac": ntt_mac_s / mac_peak, "ms_per_transform": ntt["ms"] / max(1, n_transforms),
                    "note": "the transform is IMAD-bound at 254 bits (25 IMAD/B vs a machine balance of ~3): imad_frac is the binding fraction"}
    msm_ms = stage["msm_a_ms"] + stage["msm_b1_ms"] + stage["msm_k_ms"] + stage["msm_z_ms"]
    pts_per_rank = (inst.full["a"] + inst.full["b"] + inst.full["k"] + inst.full["z"]) / world
    value = step_ms / 1e3
    line = {
        "metric": METRIC, "value": value, "unit": "s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": step_ms, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "u32 limbs (254-bit Montgomery)", "data": "synthetic",
        "config": workload_config(inst.log_n, world),
        "e2e": {"value": e2e_s, "unit": "s", "h2d_bytes_per_step": h2d_bytes,
                "d2h_bytes_per_step": int(4 * 16 * 128 + 16 * 256),
                "path": "b200_groth16_prove_core (pinned host buffers) + b200_groth16_finalize"},
        "gpu_launches": prof["kernel_launches"],
        "roofline": roofline, "roofline_ntt": roofline_ntt,
        "clocks": clk,
        "sub_metrics": {
            "g1_msm_points_per_s": pts_per_rank / (msm_ms * 1e-3) * world if msm_ms else None,
            "g2_msm_points_per_s": (inst.full["b"] / world) / (stage["msm_b2_ms"] * 1e-3) * world if stage["msm_b2_ms"] else None,
            "ntt_hbm_gbs": ntt_gbs, "stage_ms": stage,
            "kernel_ms_per_step": {k: v["ms"] / args.steps for k, v in prof.items() if isinstance(v, dict)},
            "kernel_ms_note": "msm_sort (and part of msm_reduce) runs on a second stream under the previous MSM's accumulation: "
                              "those spans overlap the others and are stretched by sharing the SMs; they do not add up to the step",
            "g2_acc_frac_of_imad_peak": (g2["units"] * MAC32_PER_G2_POINT / (g2["ms"] * 1e-3) / mac_peak) if g2["ms"] else None,
            "pk_sets": pk_info, "imad_peak": peak, "verified": verified, "verifier": "satisfying witness of the refCircuit; all five MSM results equal [e]G for the discrete logs "
            "the witness implies under the known toxic waste; verify.go:65-137 holds in the exponent; fix-ups checked (groth16_host)", "setup_s": t_setup, "wall_s_timed_region": wall_dev, "e2e_finalize_ms": getattr(args, "finalize_ms", None),
        },
    }
    if getattr(args, "plonk_result", None):
        line["plonk"] = args.plonk_result
    if getattr(args, "bls_result", None) is not None:
        line["bls12_377"] = args.bls_result
    if getattr(args, "cpu_result", None) is not None:
        line["cpu_baseline"] = args.cpu_result
    print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--log-n", type=int, default=24)
    ap.add_argument("--cpu-sample-log", type=int, default=20, help="size of the cpu_baseline sample of the GPU arm's line")
    ap.add_argument("--reference-budget", type=float, default=150.0, help="--impl reference: seconds of timed full-size CPU steps")
    ap.add_argument("--plonk-log-n", type=int, default=22, help="also time BN254 PlonK Prove at 2^k gates (config 4; sharded over the ranks at N > 1); 0 = skip")
    ap.add_argument("--bls12-377-log-n", type=int, default=22, help="N = 1: also time BLS12-377 Groth16 Prove at 2^k constraints (config 5) through libb200prover_bls12377.so; 0 = skip")
    ap.add_argument("--no-plonk-dist", dest="plonk_dist", action="store_false", help="N > 1 under torchrun: skip the sharded PlonK proof (config 4 on N GPUs)")
    ap.add_argument("--opt", action="append", default=[], metavar="KEY=VALUE", help="b200_set_option before the key is built (tuning sweeps only; the default run sets none)")
    ap.add_argument("--devices", default="", help="N > 1: device ordinals of the group, comma separated (default 0..N-1; a device may repeat: virtual ranks, functional runs on one GPU)")
    ap.add_argument("--no-check", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    return gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
