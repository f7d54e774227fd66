"""The persistent (cooperative) sort / reduction-tail kernels, the SRS slice entry point and the multi-GPU group of
the C ABI: every variant must give the same BYTES as the round-1 multi-launch path and as the oracle."""
import json
import os
import random
import subprocess
import sys

import pytest

from oracle.kzg_oracle import get_curve
from tests.golden_util import CURVE_NAMES, ROOT

sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import kzg_b200  # noqa: E402

pytestmark = pytest.mark.gpu


def _rand_bytes(C, rng, n):
    return C.scalars_to_bytes([rng.randrange(C.r) for _ in range(n)])


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_tail_and_fused_sort_give_the_same_bytes(cname):
    C = get_curve(cname)
    ctx = kzg_b200.Context(cname, 0)
    rng = random.Random(77 + C.id)
    s = rng.randrange(C.r)
    nmax = (1 << 14) + 3
    nb2 = 2 * ctx.M
    for pre in (1, 0):
        ctx.set_option("precompute", pre)
        ctx.set_option("tail", 1)
        ctx.setup_powers(s, nmax, want_bytes=False)
        # (batches of more than 592 MSMs finish in a separate launch after the tail kernel's combine step)
        for count, batch in ((1, 1), (5, 1), (300, 3), (4097, 1), (4097, 5), (nmax, 1), (2000, 40), (600, 700)):
            coeffs = [rng.randrange(C.r) for _ in range(count * batch)]
            raw = C.scalars_to_bytes(coeffs)
            ctx.set_option("tail", 1)
            got, infs = ctx.msm_g1_batch(raw, count, batch)       # projective 8-lane tail (default)
            ctx.set_option("tail_proj", 0)
            assert ctx.msm_g1_batch(raw, count, batch) == (got, infs), (pre, count, batch)  # XYZZ 4-lane tail
            ctx.set_option("tail_proj", 1)
            ctx.set_option("tail", 0)
            old, infs0 = ctx.msm_g1_batch(raw, count, batch)      # round-1 multi-launch path
            assert got == old and infs == infs0, (pre, count, batch)
            if count <= 4097:  # against the oracle through the known secret: sum c_i [s^i]G = [P(s)]G
                want = C.g1_to_octet(C.mul(C.poly_eval(coeffs[:count], s), C.g1))[1:]
                assert got[:nb2] == want, (pre, count, batch)
        # structured scalars: one bucket holds every record (deep fold inside the tail kernel), r - 1, half zeros
        ctx.set_option("tail", 1)
        for name, sc in (("ones", [1] * nmax), ("rm1", [C.r - 1] * nmax),
                         ("half_zero", [rng.randrange(C.r) if i & 1 else 0 for i in range(nmax)]),
                         ("single", [0] * (nmax - 1) + [rng.randrange(1, C.r)]), ("zeros", [0] * 100)):
            raw = C.scalars_to_bytes(sc)
            got = ctx.msm_g1(raw, len(sc))
            ctx.set_option("tail_proj", 0)
            assert ctx.msm_g1(raw, len(sc)) == got, (pre, name)
            ctx.set_option("tail_proj", 1)
            ctx.set_option("tail", 0)
            assert ctx.msm_g1(raw, len(sc)) == got, (pre, name)
            ctx.set_option("tail", 1)
            want = C.mul(C.poly_eval(sc, s), C.g1)
            assert got == (C.g1_to_octet(want)[1:], want is None), (pre, name)
    # G2 through the same kernels
    ctx.set_option("precompute", 2)
    ctx.setup_powers(s, 700)
    coeffs = [rng.randrange(C.r) for _ in range(700)]
    got = ctx.msm_g2(C.scalars_to_bytes(coeffs), 700)
    assert got[0] == C.g2_to_octet(C.mul(C.poly_eval(coeffs, s), C.g2, g2=True))[1:]
    ctx.set_option("tail", 0)
    assert ctx.msm_g2(C.scalars_to_bytes(coeffs), 700) == got
    ctx.close()


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_g2_projective_tail_gives_the_same_bytes(cname):
    """"tail_proj" = 2: the G2 reduction tail in complete projective coordinates on whole-warp teams (msm_tail_proj.cuh,
    ProjTeam<Fp2>) against the XYZZ team tail, the round-1 multi-launch path and the oracle."""
    C = get_curve(cname)
    ctx = kzg_b200.Context(cname, 0)
    rng = random.Random(99 + C.id)
    s = rng.randrange(C.r)
    nmax = 4097
    for pre in (2, 0):
        ctx.set_option("precompute", pre)
        ctx.setup_powers(s, nmax)
        cases = [("n%d" % n, [rng.randrange(C.r) for _ in range(n)]) for n in (1, 2, 3, 129, 1025, nmax)]
        cases += [("ones", [1] * 600), ("rm1", [C.r - 1] * 600), ("zeros", [0] * 50),
                  ("half_zero", [rng.randrange(C.r) if i & 1 else 0 for i in range(nmax)]),
                  ("single", [0] * 999 + [rng.randrange(1, C.r)])]
        for name, sc in cases:
            raw = C.scalars_to_bytes(sc)
            ctx.set_option("tail", 1)
            ctx.set_option("tail_proj", 2)
            got = ctx.msm_g2(raw, len(sc))
            ctx.set_option("tail_proj", 1)
            assert ctx.msm_g2(raw, len(sc)) == got, (pre, name)
            ctx.set_option("tail", 0)
            assert ctx.msm_g2(raw, len(sc)) == got, (pre, name)
            if len(sc) <= 1025:
                want = C.mul(C.poly_eval(sc, s), C.g2, g2=True)
                assert got == (C.g2_to_octet(want)[1:], want is None), (pre, name)
        # the sum is the point at infinity: c0 + c1 s = 0
        sc = [(C.r - 7 * s % C.r) % C.r, 7]
        ctx.set_option("tail", 1)
        ctx.set_option("tail_proj", 2)
        xy, inf = ctx.msm_g2(C.scalars_to_bytes(sc), 2)
        assert inf and xy == bytes(len(xy)), pre
        ctx.set_option("tail_proj", 1)
    ctx.close()


@pytest.mark.parametrize("cname", ["bls12381", "bn254"])
def test_table_ladder_gives_the_same_bytes(cname):
    """A large SRS keeps window tables of its first 8192 / 131072 points too (ctx.cuh TableRung); MSMs that fit use them.
    Same bytes as with the ladder switched off (main table), as the no-table path, and as the oracle."""
    C = get_curve(cname)
    ctx = kzg_b200.Context(cname, 0)
    ctx.set_option("setup_g2", 0)
    rng = random.Random(123 + C.id)
    s = rng.randrange(C.r)
    n = 40000                      # >= 4 * 8192: the first rung exists, the second does not
    ctx.setup_powers(s, n, want_bytes=False)
    nb2 = 2 * ctx.M
    for count, batch, off in ((1, 1, 0), (2, 1, 0), (4097, 1, 0), (8192, 1, 0), (8193, 1, 0), (4097, 3, 0), (100, 1, 8000),
                              (100, 1, 8150), (500, 2, 7000)):
        coeffs = [rng.randrange(C.r) for _ in range(count * batch)]
        raw = C.scalars_to_bytes(coeffs)
        call = (lambda: ctx.msm_g1_batch(raw, count, batch)) if batch > 1 else (lambda: ctx.msm_g1(raw, count, off))
        ctx.set_option("table_ladder", 1)
        got = call()
        s0 = ctx.msm_stats()
        ctx.set_option("table_ladder", 0)
        assert call() == got, (count, batch, off)
        s1 = ctx.msm_stats()
        if off + count <= 8192 and cname == "bls12381":
            assert s0["window_bits"] < s1["window_bits"], (count, s0, s1)   # the rung's narrower windows were used
        elif off + count <= 8192:  # (BN254: the cost model picks the same width for 8192 and 40000 points: no rung)
            assert s0["window_bits"] <= s1["window_bits"], (count, s0, s1)
        else:
            assert s0["window_bits"] == s1["window_bits"], (count, s0, s1)
        if count <= 4097 and batch == 1:
            want = C.mul(C.poly_eval(coeffs, s) * pow(s, off, C.r) % C.r, C.g1)
            assert got[0] == C.g1_to_octet(want)[1:], (count, off)
    ctx.close()


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_direct_lookup_table_gives_the_same_bytes(cname):
    """"direct_table_gb": every multiple a digit can select is tabulated (csrc/msm_direct.cuh); G1 MSMs become plain sums of
    looked-up points.  Same bytes as the bucket pipeline and as the oracle: single MSMs, batches that finish inside the tail
    kernel and larger ones, point ranges, edge scalars, a degenerate SRS (secret 1: every point equal; secret 0: infinities)."""
    C = get_curve(cname)
    rng = random.Random(321 + C.id)
    nb2 = None
    for secret, n in ((rng.randrange(2, C.r), 300), (1, 40), (0, 40)):
        ctx = kzg_b200.Context(cname, 0)
        ctx.set_option("setup_g2", 0)
        ctx.set_option("direct_table_gb", 1)     # 300 points: c = 11 or 12 fits 1 GiB
        ctx.setup_powers(secret, n, want_bytes=False)
        nb2 = 2 * ctx.M
        cases = [(1, 1, 0), (2, 1, 0), (n, 1, 0), (n - 3, 1, 3), (n, 7, 0), (17, 700, 0), (n, 40, 0)]
        for count, batch, off in cases:
            coeffs = [rng.randrange(C.r) for _ in range(count * batch)]
            if batch == 7:
                coeffs[:count] = [0] * count                      # an all-zero polynomial inside a batch
                coeffs[count:2 * count] = [C.r - 1] * count
                coeffs[2 * count] = 1
            raw = C.scalars_to_bytes(coeffs)
            call = (lambda: ctx.msm_g1_batch(raw, count, batch)) if batch > 1 else (lambda: ctx.msm_g1(raw, count, off))
            ctx.set_option("direct_table_gb", 1)
            got = call()
            ctx.set_option("direct_table_gb", 0)                  # the table stays; the bucket pipeline runs
            assert call() == got, (secret, count, batch, off)
            if batch <= 7:
                for b in range(batch):
                    want = C.mul(C.poly_eval(coeffs[b * count:(b + 1) * count], secret) * pow(secret, off, C.r) % C.r, C.g1)
                    xy = got[0][b * nb2:(b + 1) * nb2]
                    inf = got[1][b] if batch > 1 else got[1]
                    assert (xy, inf) == (C.g1_to_octet(want)[1:], want is None), (secret, count, batch, off, b)
        # the sum is the point at infinity: c0 + c1 s = 0
        if secret > 1:
            ctx.set_option("direct_table_gb", 1)
            sc = [(C.r - 9 * secret % C.r) % C.r, 9]
            xy, inf = ctx.msm_g1(C.scalars_to_bytes(sc), 2)
            assert inf and xy == bytes(len(xy))
        ctx.close()


def test_tail_at_2_16_and_2_18():
    """sizes where part of the reduction is still wide launches and the rest runs in the tail kernel"""
    C = get_curve("bls12381")
    ctx = kzg_b200.Context("bls12381", 0)
    ctx.set_option("setup_g2", 0)
    rng = random.Random(5)
    s = rng.randrange(C.r)
    for lg in (16, 18):
        n = 1 << lg
        ctx.setup_powers(s, n, want_bytes=False)
        stats = None
        raw = os.urandom(n * 32)
        raw = bytes(b & 0x3F if (i & 31) == 31 else b for i, b in enumerate(raw)) if lg == 16 else \
            b"".join(os.urandom(31) + b"\x11" for _ in range(n))
        ctx.set_option("tail", 1)
        got = ctx.msm_g1(raw, n)
        l0 = ctx.launch_count
        ctx.msm_g1(raw, n)
        launches = ctx.launch_count - l0
        ctx.set_option("tail", 0)
        assert ctx.msm_g1(raw, n) == got, lg
        assert launches <= 25, (lg, launches)
        if lg == 16:
            sc = C.scalars_from_bytes(raw)
            assert got[0] == C.g1_to_octet(C.mul(C.poly_eval(sc, s), C.g1))[1:]
    ctx.close()


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_setup_powers_range_matches_oracle(cname):
    C = get_curve(cname)
    ctx = kzg_b200.Context(cname, 0)
    s = random.Random(11).randrange(C.r)
    first, n = 1000003, 7
    g1b, g2b = ctx.setup_powers(s, n, first_index=first)
    for i in range(n):
        e = pow(s, first + i, C.r)
        assert g1b[i * 2 * ctx.M:(i + 1) * 2 * ctx.M] == C.g1_to_octet(C.mul(e, C.g1))[1:], i
        assert g2b[i * 4 * ctx.M:(i + 1) * 4 * ctx.M] == C.g2_to_octet(C.mul(e, C.g2, g2=True))[1:], i
    # the slice is resident as points 0..n-1
    sc = [random.Random(i).randrange(C.r) for i in range(n)]
    xy, inf = ctx.msm_g1(C.scalars_to_bytes(sc), n)
    e = sum(c * pow(s, first + i, C.r) for i, c in enumerate(sc)) % C.r
    assert xy == C.g1_to_octet(C.mul(e, C.g1))[1:]
    ctx.close()


def _group_devices():
    lib = kzg_b200.load_library()
    nd = lib.kzg_gpu_device_count()
    return [0, 1] if nd >= 2 else [0, 0]  # one GPU: two members sharing it (same code paths, no speed-up)


def test_group_blob_sharding_and_point_range_split():
    cname = "bls12381"
    C = get_curve(cname)
    rng = random.Random(21)
    s = rng.randrange(C.r)
    single = kzg_b200.Context(cname, 0)
    n = 4096 + 17
    g1b, g2b = single.setup_powers(s, n)
    for devices, loader in ((_group_devices(), "setup"), ([0, 0, 0], "load"), ([0], "load")):
        grp = kzg_b200.Group(cname, devices)
        assert grp.size == len(devices)
        grp.set_option("split_min", 1024)
        if loader == "setup":
            a, b = grp.setup_powers(s, n)
            assert a == g1b and b == g2b
        else:
            grp.srs_load(g1b, g2b)
        # one MSM split by point range == the single-device bytes, for whole and partial ranges
        sc = _rand_bytes(C, rng, n)
        for count, off in ((n, 0), (3000, 500), (1024, 3000), (100, 0), (0, 0)):
            want = single.msm_g1(sc[off * 32:(off + count) * 32], count, base_offset=off) if count else (bytes(96), True)
            assert grp.msm_g1(sc[off * 32:(off + count) * 32] if count else b"", count, base_offset=off) == want, (devices, count, off)
        # blob batch sharded by blob
        for count, batch in ((1000, 37), (4096, 3), (7, 1)):
            raw = _rand_bytes(C, rng, count * batch)
            assert grp.msm_g1_batch(raw, count, batch) == single.msm_g1_batch(raw, count, batch), (devices, count, batch)
        raw = _rand_bytes(C, rng, 2048 * 9)
        roots = [rng.randrange(0, 3000) for _ in range(9)]
        assert grp.prove_batch(raw, 2048, roots, 3) == single.prove_batch(raw, 2048, roots, 3)
        with pytest.raises(RuntimeError):
            grp.msm_g1(sc, n + 1)
        grp.close()
    single.close()


def test_two_ranks_nccl_fold_matches_oracle():
    """One process per GPU over NCCL (the torchrun form of the point-range split and of the blob sharding): bench.py
    asserts the folded bytes against [P(s)]G1 on every path and exits non-zero otherwise."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 29600 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "2", "--warmup", "3",
           "--log2n", "16", "--strong-log2n", "17", "--blobs", "65", "--no-cpu"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stderr[-3000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["n_gpus"] == 2 and line["parity_asserted"] is True
    assert line["strong"]["msm_2p17"]["parity_asserted"] and line["commit"]["parity_asserted"]
    assert line["commit"]["blobs"] == 65


def test_quotient_paths_for_long_polynomials():
    """Large-degree polynomials (the reference's --benchmark-common path, benchmark/benchmark.cpp:123-136): k rounds over
    the whole GPU for long polynomials with few roots, the power-series path beyond the round-1 cap of 16384 quotient
    coefficients; checked against the oracle's synthetic division and against the independent systolic kernel."""
    cname = "bn254"
    C = get_curve(cname)
    ctx = kzg_b200.Context(cname, 0)
    rng = random.Random(31)
    n = (1 << 17) + 3
    P = [rng.randrange(C.r) for _ in range(n)]
    raw = C.scalars_to_bytes(P)
    for off, k in ((5, 1), (123456789, 3)):
        q = list(P)
        for j in range(k):  # k synthetic divisions by (x - (off + j))
            a = (off + j) % C.r
            out = [0] * (len(q) - 1)
            carry = 0
            for i in range(len(q) - 1, 0, -1):
                carry = (q[i] + carry * a) % C.r
                out[i - 1] = carry
            q = out
        assert ctx.poly_quotient(raw, off, k) == C.scalars_to_bytes(q), (off, k)
    # proof over the long polynomial: quotient (grid rounds) + MSM of 2^17 + 2 terms
    s = rng.randrange(C.r)
    ctx.set_option("setup_g2", 0)
    ctx.setup_powers(s, n, want_bytes=False)
    xy, inf = ctx.prove(raw, 5, 1)
    qs = (C.poly_eval(P, s) - C.poly_eval(P, 5)) * pow(s - 5, -1, C.r) % C.r
    assert xy == C.g1_to_octet(C.mul(qs, C.g1))[1:]
    # many roots, quotient longer than 16384 coefficients: series path == systolic path
    n2, k2 = 40000, 18000
    raw2 = C.scalars_to_bytes([rng.randrange(C.r) for _ in range(n2)])
    q_series = ctx.poly_quotient(raw2, 7, k2)
    ctx.set_option("long_division", 0)
    assert ctx.poly_quotient(raw2, 7, k2) == q_series
    ctx.set_option("long_division", 3)
    ctx.close()
