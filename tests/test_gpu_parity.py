"""Parity tests proper: the CUDA path, called through the C ABI (include/kzg_gpu.h) via the Python mirror
of kzg.h, against (a) the golden vectors produced by the UNMODIFIED reference (tests/golden, bit-exact on
commit.serialize(), proof.serialize(), polyeval_G2, quotient and evaluation bytes), (b) the CPU oracle on
seeded inputs, (c) size-independent properties at full size, and (d) the reference's own verify_proof
(oracle/_ref/kzgref-<curve>, prebuilt) accepting GPU proofs and refuting tampered data.
Everything here is exact integer arithmetic: the bar is byte equality."""
import os
import random
import subprocess
import sys
import tempfile

import pytest

from oracle.kzg_oracle import get_curve
from tests.golden_util import CURVE_NAMES, ROOT, load

sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import kzg_b200  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctxs():
    made = {}

    def get(cname):
        if cname not in made:
            made[cname] = kzg_b200.Context(cname, 0)
        return made[cname]

    yield get
    for c in made.values():
        c.close()


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_device_selftest(ctxs, cname):
    assert ctxs(cname).selftest() == 0


@pytest.mark.parametrize("cname", CURVE_NAMES)
@pytest.mark.parametrize("precompute", [0, 1])
def test_golden_vectors_bit_exact(ctxs, cname, precompute):
    C = get_curve(cname)
    ctx = ctxs(cname)
    kat, srs = load(cname)
    ctx.set_option("precompute", precompute)
    ctx.set_option("window_bits", 7 if precompute else 0)
    ts = kzg_b200.TrustedSetup.from_file_bytes(cname, srs, ctx=ctx)
    assert ts.export_setup() == srs
    for case in kat["cases"]:
        P = kzg_b200.Poly(ctx, bytes.fromhex(case["coeffs"]))
        assert P.serialize().hex() == case["polyser"], case["name"]
        assert kzg_b200.Poly.deserialize((ctx.SB, ctx.r), P.serialize()).coeffs == P.coeffs
        commit = ts.create_commit(P)
        assert commit.serialize().hex() == case["commit"], case["name"]
        assert ts.verify_commit(kzg_b200.Commit.deserialize(commit.serialize()), P)
        if len(P):
            xy, inf = ctx.msm_g2(P.raw, len(P))
            assert (b"\x04" + xy).hex() == case["g2msm"][8:], case["name"]
        for op in case["openings"]:
            off, k = op["off"], op["k"]
            proof = ts.create_proof(P, off, k)
            assert proof.serialize().hex() == op["proof"], (case["name"], off, k)
            if len(P):
                assert ts.evaluate(P, off, k).hex() == op["ys"], (case["name"], off, k)
            if k < len(P):
                assert ctx.poly_quotient(P.raw, off, k).hex() == op["quotient"], (case["name"], off, k)
            # the verifier-side MSMs on the reference's own Z and I
            Z = kzg_b200.Poly(ctx, bytes.fromhex(op["vanishing"]))
            I = kzg_b200.Poly(ctx, bytes.fromhex(op["interp"]))
            (zxy, zinf), (ixy, iinf) = ts.verify_proof_msms(Z, I)
            g1, g2 = C.srs_from_file_bytes(srs)
            assert zxy == C.g2_to_octet(C.msm(Z.coeffs, g2, g2=True))[1:]
            assert ixy == C.g1_to_octet(C.msm(I.coeffs, g1))[1:]
    ctx.set_option("precompute", 0)
    ctx.set_option("window_bits", 0)


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_setup_powers_matches_oracle(ctxs, cname):
    C = get_curve(cname)
    ctx = ctxs(cname)
    s = random.Random(5).randrange(C.r)
    ts = kzg_b200.TrustedSetup.generate(cname, 9, s, ctx=ctx)
    g1, g2 = C.setup(s, 9)
    assert ts.export_setup() == C.srs_to_file_bytes(g1, g2)
    with pytest.raises(ValueError):
        kzg_b200.TrustedSetup.generate(cname, 1, s, ctx=ctx)


def _reference_verify(cname, srs_bytes, commit, proof, off, ys):
    binary = os.path.join(ROOT, "oracle", "_ref", f"kzgref-{cname}")
    if not os.path.exists(binary):
        pytest.skip("oracle/_ref not built")
    with tempfile.TemporaryDirectory() as td:
        paths = {}
        for name, data in (("srs", srs_bytes), ("cm", commit), ("pf", proof), ("ys", ys)):
            paths[name] = os.path.join(td, name)
            open(paths[name], "wb").write(data)
        out = subprocess.run([binary, "verify", paths["srs"], paths["cm"], paths["pf"], str(off), paths["ys"]],
                             check=True, capture_output=True, text=True).stdout
        return int(out.strip())


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_reference_verify_accepts_gpu_proofs(ctxs, cname):
    """GPU-generated SRS, commit, proof and opened values go through the reference's verify_proof."""
    C = get_curve(cname)
    ctx = ctxs(cname)
    rng = random.Random(17 + C.id)
    ts = kzg_b200.TrustedSetup.generate(cname, 300, rng.randrange(C.r), ctx=ctx)
    srs_bytes = ts.export_setup()
    P = kzg_b200.Poly(ctx, [rng.randrange(C.r) for _ in range(257)])
    commit = ts.create_commit(P)
    for off, k in ((0, 1), (5, 1), (3, 8), (100, 57)):
        proof = ts.create_proof(P, off, k)
        ys = ts.evaluate(P, off, k)
        assert ys == C.scalars_to_bytes(C.poly_eval_range(P.coeffs, off, k))
        assert _reference_verify(cname, srs_bytes, commit.serialize(), proof.serialize(), off, ys) == 1
        bad = bytearray(ys)
        bad[0] ^= 1
        assert _reference_verify(cname, srs_bytes, commit.serialize(), proof.serialize(), off, bytes(bad)) == 0
        assert _reference_verify(cname, srs_bytes, commit.serialize(), proof.serialize(), off + 1, ys) == 0


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_msm_degree_4096_vs_oracle_and_options(ctxs, cname):
    """BASELINE configs 1/2: a degree-4096 commitment.  Oracle check through the known secret:
    sum c_i [s^i]G = [P(s)]G, one oracle scalar multiplication (size-independent)."""
    C = get_curve(cname)
    ctx = ctxs(cname)
    rng = random.Random(4096 + C.id)
    s = rng.randrange(C.r)
    n = 4200
    ts = kzg_b200.TrustedSetup.generate(cname, n, s, ctx=ctx)
    coeffs = [rng.randrange(C.r) for _ in range(4097)]
    P = kzg_b200.Poly(ctx, coeffs)
    want = C.serialize_g1(C.mul(C.poly_eval(coeffs, s), C.g1))
    assert ts.create_commit(P).serialize() == want
    # every window size / task split / table mode gives the same bytes
    for wb, pre, chunk in ((8, 0, 7), (11, 0, 64), (13, 1, 16), (10, 1, 1000)):
        ctx.set_option("window_bits", wb)
        ctx.set_option("precompute", pre)
        ctx.set_option("task_chunk", chunk)
        ctx.srs_load(ts.g1_bytes, ts.g2_bytes)
        assert ts.create_commit(P).serialize() == want, (wb, pre, chunk)
    ctx.set_option("window_bits", 0)
    ctx.set_option("precompute", 0)
    ctx.set_option("task_chunk", 0)
    ctx.srs_load(ts.g1_bytes, ts.g2_bytes)
    # single-point proof: q(s) = (P(s) - P(a)) / (s - a)
    a = 7
    qs = (C.poly_eval(coeffs, s) - C.poly_eval(coeffs, a)) * pow(s - a, -1, C.r) % C.r
    assert ts.create_proof(P, a, 1).serialize() == C.serialize_g1(C.mul(qs, C.g1))
    # multi-point proof (k = 64) through the vanishing polynomial
    k, off = 64, 11
    zs = 1
    for j in range(k):
        zs = zs * (s - off - j) % C.r
    q, I, _ = C.poly_quotient(coeffs, off, k)
    assert ts.create_proof(P, off, k).serialize() == C.serialize_g1(C.mul(C.poly_eval(q, s), C.g1))
    assert ctx.poly_quotient(P.raw, off, k) == C.scalars_to_bytes(q)
    # G2 MSM over 4097 terms (verify_proof's polyeval_G2 at k = 4096)
    xy, inf = ctx.msm_g2(P.raw, len(P))
    assert xy == C.g2_to_octet(C.mul(C.poly_eval(coeffs, s), C.g2, g2=True))[1:]
    # error behaviour of create_commit / create_proof (src/trusted_setup.cpp:138-139, 215-216)
    with pytest.raises(ValueError):
        ts.create_commit([1] * n)
    with pytest.raises(ValueError):
        ts.create_proof(P, 0, 0)
    assert ts.create_proof(P, 0, 4097).is_inf and ts.create_proof(P, 0, 5000).is_inf


def test_edge_scalars_and_structured_inputs(ctxs):
    cname = "bls12381"
    C = get_curve(cname)
    ctx = ctxs(cname)
    rng = random.Random(3)
    s = rng.randrange(C.r)
    n = 600
    ts = kzg_b200.TrustedSetup.generate(cname, n + 1, s, ctx=ctx)
    cases = {
        "zeros": [0] * n,
        "ones": [1] * n,
        "rm1": [C.r - 1] * n,
        "single": [0] * (n - 1) + [rng.randrange(1, C.r)],
        "half_zero": [rng.randrange(C.r) if i % 2 else 0 for i in range(n)],
        "small": [rng.randrange(256) for _ in range(n)],
        "pow2": [1 << (i % 255) for i in range(n)],
    }
    for name, sc in cases.items():
        raw = C.scalars_to_bytes(sc)
        xy, inf = ctx.msm_g1(raw, n)
        want = C.mul(C.poly_eval(sc, s), C.g1) if any(sc) else None
        assert xy == C.g1_to_octet(want)[1:], name
        assert inf == (want is None), name
    # non-canonical scalars (>= r) are reduced like PAIR_G1mul's BIG_mod
    sc = [(1 << 256) - 1 - i for i in range(40)]
    raw = b"".join(v.to_bytes(32, "little") for v in sc)
    xy, _ = ctx.msm_g1(raw, 40)
    assert xy == C.g1_to_octet(C.mul(C.poly_eval([v % C.r for v in sc], s), C.g1))[1:]
    # a degenerate SRS (secret 1: every base is the generator) drives the doubling / cancellation branches
    ts1 = kzg_b200.TrustedSetup.generate(cname, 64, 1, ctx=ctx)
    sc = [5] * 20 + [C.r - 5] * 20 + [3] * 10
    xy, inf = ctx.msm_g1(C.scalars_to_bytes(sc), len(sc))
    assert xy == C.g1_to_octet(C.mul(30, C.g1))[1:]
    xy, inf = ctx.msm_g1(C.scalars_to_bytes([9] * 8 + [C.r - 9] * 8), 16)
    assert inf and xy == bytes(96)
    # base_offset, empty input
    ts = kzg_b200.TrustedSetup.generate(cname, 50, s, ctx=ctx)
    sc = [rng.randrange(C.r) for _ in range(10)]
    xy, _ = ctx.msm_g1(C.scalars_to_bytes(sc), 10, base_offset=7)
    assert xy == C.g1_to_octet(C.mul(C.poly_eval(sc, s) * pow(s, 7, C.r), C.g1))[1:]
    xy, inf = ctx.msm_g1(b"", 0)
    assert inf and xy == bytes(96)
    with pytest.raises(RuntimeError):
        ctx.msm_g1(C.scalars_to_bytes(sc), 10, base_offset=45)


def test_batch_and_point_range_split(ctxs):
    """BASELINE config 3 (batched blob commitments) and config 4 (point-range split + partial-sum fold)."""
    cname = "bls12381"
    C = get_curve(cname)
    ctx = ctxs(cname)
    rng = random.Random(8)
    s = rng.randrange(C.r)
    n, batch = 1024, 6
    ts = kzg_b200.TrustedSetup.generate(cname, n, s, ctx=ctx)
    blobs = [[rng.randrange(C.r) for _ in range(n)] for _ in range(batch)]
    blobs[2] = [0] * n
    raw = b"".join(C.scalars_to_bytes(b) for b in blobs)
    out, infs = ctx.msm_g1_batch(raw, n, batch)
    for b in range(batch):
        want = C.mul(C.poly_eval(blobs[b], s), C.g1) if any(blobs[b]) else None
        assert out[b * 96:(b + 1) * 96] == C.g1_to_octet(want)[1:], b
        assert infs[b] == (want is None)
    # split one MSM by point range into 4 shards, fold the projective partial sums
    full, _ = ctx.msm_g1(C.scalars_to_bytes(blobs[0]), n)
    partials = []
    for g in range(4):
        lo, hi = g * n // 4, (g + 1) * n // 4
        ctx.msm_g1(C.scalars_to_bytes(blobs[0][lo:hi]), hi - lo, base_offset=lo)
        partials.append(ctx.msm_fetch_partial())
    xy, inf = ctx.g1_sum_partials(partials)
    assert xy == full and not inf
    # the same fold with the partials lying in device memory (what the NCCL all-gather leaves behind)
    import torch
    import sharding
    dev = []
    for g in range(4):
        lo, hi = g * n // 4, (g + 1) * n // 4
        ctx.msm_g1(C.scalars_to_bytes(blobs[0][lo:hi]), hi - lo, base_offset=lo)
        ptr, nb = ctx.msm_partial_device()
        assert nb == len(partials[g])
        dev.append(torch.as_tensor(sharding._DevBytes(ptr, nb), device="cuda").clone())
    allp = torch.cat(dev)
    torch.cuda.synchronize()
    xy, inf = ctx.g1_sum_partials_device(allp, 4)
    assert xy == full and not inf
    # ranks of a point-range split skip their own affine normalisation (finish_affine = 0): the projective partial is
    # the same group element (its representation depends on the order the atomics sorted the records in); and a
    # fold of many partials (more than the 128 teams of the fold kernel hold at once, infinity included)
    ctx.set_option("finish_affine", 0)
    ctx.msm_g1(C.scalars_to_bytes(blobs[0][:n // 4]), n // 4, base_offset=0)
    first = ctx.msm_fetch_partial()
    ctx.set_option("finish_affine", 1)
    xy, inf = ctx.g1_sum_partials([first] + partials[1:])
    assert xy == full and not inf
    many = []
    for g in range(150):
        lo, hi = g * n // 150, (g + 1) * n // 150
        ctx.msm_g1(C.scalars_to_bytes(blobs[0][lo:hi]), hi - lo, base_offset=lo)
        many.append(ctx.msm_fetch_partial())
    many.insert(7, bytes(len(many[0])))  # the point at infinity (ZZ = 0)
    xy, inf = ctx.g1_sum_partials(many)
    assert xy == full and not inf


def test_msm_2_16_linearity_and_oracle(ctxs):
    """Config 4 at 2^16: [P(s)]G check plus linearity MSM(a) + MSM(b) = MSM(a+b) at full size."""
    cname = "bls12381"
    C = get_curve(cname)
    ctx = ctxs(cname)
    rng = random.Random(16)
    s = rng.randrange(C.r)
    n = 1 << 16
    ctx.setup_powers(s, n, want_bytes=False)
    a = [rng.randrange(C.r) for _ in range(n)]
    b = [rng.randrange(C.r) for _ in range(n)]
    ab = [(x + y) % C.r for x, y in zip(a, b)]

    def ps(v):
        acc, si = 0, 1
        for c in v:
            acc = (acc + c * si) % C.r
            si = si * s % C.r
        return acc

    xa, _ = ctx.msm_g1(C.scalars_to_bytes(a), n)
    assert xa == C.g1_to_octet(C.mul(ps(a), C.g1))[1:]
    xab, _ = ctx.msm_g1(C.scalars_to_bytes(ab), n)
    xb, _ = ctx.msm_g1(C.scalars_to_bytes(b), n)
    A = (int.from_bytes(xa[:48], "big"), int.from_bytes(xa[48:], "big"))
    B = (int.from_bytes(xb[:48], "big"), int.from_bytes(xb[48:], "big"))
    assert xab == C.g1_to_octet(C.add(A, B))[1:]
    ctx.set_option("precompute", 1)
    ctx.setup_powers(s, n, want_bytes=False)
    xa2, _ = ctx.msm_g1(C.scalars_to_bytes(a), n)
    ctx.set_option("precompute", 0)
    assert xa2 == xa


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_quotient_many_roots_systolic(ctxs, cname):
    """A6 with many opened points (multi-proof rows of benchmark.cpp), against the oracle: the CTA-wide long
    division by Z on an 8-CTA cluster (default for one polynomial, option long_division = 2) and on one CTA
    (= 1; 1, 2 and 4 window elements per thread), and the systolic kernel (= 0; one pass k = 64, 512, chained passes k = 700, 1300), plus the boundary with the
    scan kernel (k = 47, 48)."""
    C = get_curve(cname)
    ctx = ctxs(cname)
    rng = random.Random(77 + C.id)
    n = 1501
    coeffs = [rng.randrange(C.r) for _ in range(n)]
    raw = C.scalars_to_bytes(coeffs)
    for long_division in (3, 2, 1, 0):
        ctx.set_option("long_division", long_division)
        for off, k in ((2, 7), (4, 8), (0, 47), (0, 48), (3, 64), (10, 512), (1, 700), (5, 1300), (0, 1500)):
            q, _, _ = C.poly_quotient(coeffs, off, k)
            assert ctx.poly_quotient(raw, off, k) == C.scalars_to_bytes(q), (long_division, off, k)
    n2 = 2600  # 2 and 4 window elements per thread (k > 1024, k > 2048), window not a multiple of the CTA size
    coeffs2 = [rng.randrange(C.r) for _ in range(n2)]
    for off, k in ((2, 1025), (0, 2048), (7, 2100)):
        q, _, _ = C.poly_quotient(coeffs2, off, k)
        for long_division in (1, 2, 3):
            ctx.set_option("long_division", long_division)
            assert ctx.poly_quotient(C.scalars_to_bytes(coeffs2), off, k) == C.scalars_to_bytes(q), (long_division, off, k)
    # proof through the chained path equals [q(s)]G
    s = rng.randrange(C.r)
    ts = kzg_b200.TrustedSetup.generate(cname, n + 1, s, ctx=ctx)
    P = kzg_b200.Poly(ctx, coeffs)
    q, _, _ = C.poly_quotient(coeffs, 2, 600)
    assert ts.create_proof(P, 2, 600).serialize() == C.serialize_g1(C.mul(C.poly_eval(q, s), C.g1))


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_batched_affine_levels(ctxs, cname):
    """Forced batched-affine pre-reduction (msm_affine.cuh) at 1-4 levels: same bytes as the XYZZ-only path and
    as the oracle, including a degenerate SRS (every base equal: P + P and P + (-P) inside the affine levels),
    zero scalars and a batch."""
    C = get_curve(cname)
    ctx = ctxs(cname)
    rng = random.Random(31 + C.id)
    s = rng.randrange(C.r)
    n = 3000
    try:
        for pre in (0, 2):
            ctx.set_option("precompute", pre)
            ctx.set_option("affine_levels", 0)
            ctx.setup_powers(s, n, want_bytes=False)
            sc = [rng.randrange(C.r) for _ in range(n)]
            sc[5] = 0
            raw = C.scalars_to_bytes(sc)
            want = C.g1_to_octet(C.mul(C.poly_eval(sc, s), C.g1))[1:]
            for lv in (0, 1, 2, 3, 4):
                ctx.set_option("affine_levels", lv)
                xy, inf = ctx.msm_g1(raw, n)
                assert xy == want and not inf, (pre, lv)
            ctx.set_option("affine_levels", 2)
            out, infs = ctx.msm_g1_batch(raw[:2 * 1000 * ctx.SB], 1000, 2)
            for b in range(2):
                part = sc[b * 1000:(b + 1) * 1000]
                assert out[b * 2 * ctx.M:(b + 1) * 2 * ctx.M] == C.g1_to_octet(C.mul(C.poly_eval(part, s), C.g1))[1:]
        # degenerate SRS (s = 1): all bases are the generator
        ctx.set_option("precompute", 0)
        ctx.setup_powers(1, 64, want_bytes=False)
        for lv in (1, 2, 3):
            ctx.set_option("affine_levels", lv)
            ctx.set_option("window_bits", 4)
            sc = [5] * 20 + [C.r - 5] * 20 + [3] * 10
            xy, inf = ctx.msm_g1(C.scalars_to_bytes(sc), len(sc))
            assert xy == C.g1_to_octet(C.mul(30, C.g1))[1:], lv
            xy, inf = ctx.msm_g1(C.scalars_to_bytes([9] * 8 + [C.r - 9] * 8), 16)
            assert inf and xy == bytes(2 * ctx.M), lv
            xy, inf = ctx.msm_g1(C.scalars_to_bytes([7] * 33), 33)
            assert xy == C.g1_to_octet(C.mul(7 * 33, C.g1))[1:], lv
    finally:
        ctx.set_option("affine_levels", -1)
        ctx.set_option("window_bits", 0)
        ctx.set_option("precompute", 2)


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_interpolation_matches_reference_and_round_trips(ctxs, cname):
    """kzg_gpu_interpolate_range against the reference's linear_roots_and_polyfit outputs stored
    in the golden vectors (I and Z for every opening), against the oracle's Lagrange interpolation, and — at the
    sizes the benchmark uses — through the round trip  evaluate(interpolate(ys)) == ys  and  Z(x_i) == 0."""
    C = get_curve(cname)
    ctx = ctxs(cname)
    kat, _ = load(cname)
    checked = 0
    for case in kat["cases"]:
        for op in case["openings"]:
            ys = bytes.fromhex(op["ys"])
            I, Z = ctx.interpolate_range(ys, op["off"])
            assert I.hex() == op["interp"], (case["name"], op["off"], op["k"])
            assert Z.hex() == op["vanishing"], (case["name"], op["off"], op["k"])
            checked += 1
    assert checked > 50
    rng = random.Random(11 + C.id)
    for off, k in ((0, 1), (7, 2), (0, 3), (5, 33), (100, 64)):
        ys = [rng.randrange(C.r) for _ in range(k)]
        I, Z = ctx.interpolate_range(C.scalars_to_bytes(ys), off)
        assert I == C.scalars_to_bytes((C.interpolate([off + i for i in range(k)], ys) + [0] * k)[:k]), (off, k)
        assert Z == C.scalars_to_bytes(C.vanishing_poly(off, k)), (off, k)
    for off, k in ((0, 4096), (3, 4229), (1, 1000)):
        ys = C.scalars_to_bytes([rng.randrange(C.r) for _ in range(k)])
        I, Z = ctx.interpolate_range(ys, off)
        assert ctx.poly_eval_range(I, off, k) == ys, (off, k)
        assert ctx.poly_eval_range(Z, off, k) == bytes(len(ys)), (off, k)
        assert Z[-ctx.SB:] == (1).to_bytes(ctx.SB, "little")
    P = kzg_b200.Poly.from_blob(ctx, [ord(ch) for ch in "hello there my name is bob"])
    assert P.coeffs == C.poly_trim(C.interpolate(list(range(26)), [ord(ch) for ch in "hello there my name is bob"]))


@pytest.mark.parametrize("cname", ["bls12381", "bn254"])
def test_prove_batch_equals_single_proofs(ctxs, cname):
    """kzg_gpu_prove_batch (one CTA per quotient, batched MSM) against kzg_gpu_prove, scan and systolic paths."""
    C = get_curve(cname)
    ctx = ctxs(cname)
    rng = random.Random(55 + C.id)
    n, batch = 700, 5
    ts = kzg_b200.TrustedSetup.generate(cname, n + 1, rng.randrange(C.r), ctx=ctx)
    polys = [[rng.randrange(C.r) for _ in range(n)] for _ in range(batch)]
    polys[3] = [0] * n
    raw = b"".join(C.scalars_to_bytes(p) for p in polys)
    for k in (1, 3, 64):
        roots = [rng.randrange(0, 50) for _ in range(batch)]
        out, infs = ctx.prove_batch(raw, n, roots, k)
        for b in range(batch):
            xy, inf = ctx.prove(C.scalars_to_bytes(polys[b]), roots[b], k)
            assert out[b * 2 * ctx.M:(b + 1) * 2 * ctx.M] == xy and infs[b] == inf, (k, b)
        assert infs[3]
    out, infs = ctx.prove_batch(raw, n, [0] * batch, n)
    assert all(infs) and out == bytes(len(out))


@pytest.mark.parametrize("cname", ["bls12381", "bn254"])
def test_prove_batch_many_roots_shared_positions(ctxs, cname):
    """A batch opened at many points but few distinct positions: Z and 1 / rev(Z) once per position, one batched truncated
    product for the quotients (ctx.cuh enqueue_quotient, batch_series) — against the systolic kernel (long_division = 0)
    and against kzg_gpu_prove one polynomial at a time."""
    C = get_curve(cname)
    ctx = ctxs(cname)
    rng = random.Random(77 + C.id)
    n, batch, k = 700, 6, 400
    ts = kzg_b200.TrustedSetup.generate(cname, n + 1, rng.randrange(C.r), ctx=ctx)
    polys = [[rng.randrange(C.r) for _ in range(n)] for _ in range(batch)]
    polys[2] = [0] * n
    polys[4] = polys[4][:500] + [0] * 200   # lower degree than the rest
    raw = b"".join(C.scalars_to_bytes(p) for p in polys)
    for roots in ([7] * batch, [7, 7, 9, 9, 7, 9], [0, 1, 2, 3, 4, 5]):
        ctx.set_option("long_division", 3)
        out, infs = ctx.prove_batch(raw, n, roots, k)
        ctx.set_option("long_division", 0)
        out0, infs0 = ctx.prove_batch(raw, n, roots, k)
        ctx.set_option("long_division", 3)
        assert out == out0 and infs == infs0, roots
        for b in range(batch):
            xy, inf = ctx.prove(C.scalars_to_bytes(polys[b]), roots[b], k)
            assert out[b * 2 * ctx.M:(b + 1) * 2 * ctx.M] == xy and infs[b] == inf, (roots, b)
        assert infs[2]
    # and against the oracle: proof = [q(s)]G1 with q from the CPU restatement of NTL's div
    out, infs = ctx.prove_batch(raw, n, [7] * batch, k)
    q, _, _ = C.poly_quotient(polys[1], 7, k)
    xy, inf = ctx.msm_g1(C.scalars_to_bytes(q), len(q))
    assert out[2 * ctx.M:4 * ctx.M] == xy and not inf


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_commit_bytes_device_chain(ctxs, cname):
    """blob::from_bytes -> poly::from_blob -> create_commit chained on the device
    (kzg_gpu_commit_bytes) against the oracle's blob encoding, interpolation and commitment, against the
    step-by-step path (Poly.from_blob + create_commit), and the reference's argument errors (src/blob.cpp:20-25)."""
    C = get_curve(cname)
    ctx = ctxs(cname)
    ctx.set_option("precompute", 2)
    ctx.set_option("window_bits", 0)
    rng = random.Random(31 + C.id)
    s = rng.randrange(1, C.r)
    ts = kzg_b200.TrustedSetup.generate(cname, 5000, s, ctx=ctx)
    for chunk, k, off in ((ctx.SB - 1, 40, 0), (1, 26, 0), (7, 33, 2), (ctx.SB - 1, 1, 5)):
        data = bytes(rng.randrange(256) for _ in range(chunk * k))
        commit, P = ts.commit_bytes(data, chunk, byte_offset=off * chunk)
        blob = C.blob_from_bytes(data, off * chunk, len(data), chunk)
        want = C.poly_trim(C.interpolate([x for x, _ in blob], [y for _, y in blob]))
        assert P.coeffs == want, (chunk, k, off)
        assert commit.serialize() == C.serialize_g1(C.mul(C.poly_eval(want, s), C.g1)), (chunk, k, off)
        assert ts.create_commit(kzg_b200.Poly.from_blob(ctx, [y for _, y in blob], off)) == commit
    # the demo's document shape: 4583 chunks of MAX_CHUNK_BYTES
    chunk = ctx.SB - 1
    data = bytes(rng.randrange(32, 127) for _ in range(chunk * 4583))
    commit, P = ts.commit_bytes(data, chunk)
    ys = [int.from_bytes(data[i * chunk:(i + 1) * chunk], "little") for i in range(4583)]
    assert ts.create_commit(kzg_b200.Poly.from_blob(ctx, ys, 0)) == commit
    assert ctx.poly_eval_range(P.raw + bytes((4583 - len(P)) * ctx.SB), 0, 4583) == C.scalars_to_bytes(ys)
    q, _, _ = C.poly_quotient(P.coeffs, 100, 4)
    assert ts.create_proof(P, 100, 4).serialize() == C.serialize_g1(C.mul(C.poly_eval(q, s), C.g1))
    with pytest.raises(ValueError):
        ts.commit_bytes(b"x" * ctx.SB, ctx.SB)
    with pytest.raises(ValueError):
        ts.commit_bytes(b"x" * 10, 3)
    with pytest.raises(ValueError):
        ts.commit_bytes(b"x" * 10, 5, byte_offset=3)
    c0, p0 = ts.commit_bytes(b"", 4)
    assert c0.is_inf and len(p0) == 0


def test_batch_host_scalars_pipelined_passes(ctxs):
    """A host-scalar batch large enough for the pipelined path of kzg_gpu_msm_g1_batch (passes of >= 128 MSMs, the
    copy of pass i + 1 under the kernels of pass i, two staging buffers): 300 blobs x 4096 coefficients in three
    ragged passes must give the bytes of the device-resident batch and of single-MSM calls, twice in a row."""
    import numpy as np
    import torch
    cname = "bls12381"
    C = get_curve(cname)
    ctx = ctxs(cname)
    ctx.set_option("precompute", 2)
    ctx.set_option("window_bits", 0)
    n, batch = 4096, 300
    ctx.set_option("setup_g2", 0)
    ctx.setup_powers(0x1234567, n + 1, want_bytes=False)
    rng = np.random.default_rng(99)
    a = rng.integers(0, 256, size=(batch * n, 32), dtype=np.uint8)
    a[:, 31] &= 0x3F  # below r
    a[5 * n:6 * n] = 0  # one all-zero blob -> infinity
    host = a.reshape(-1).tobytes()
    dev = torch.from_numpy(a.reshape(-1)).cuda()
    ctx.msm_g1_launch(dev, n, batch, 0)
    want, want_inf = ctx.msm_fetch(batch)
    # one large host MSM first (its chunked copy creates the copy stream), then the pipelined batch on the same context
    big = host[:70 * n * 32]
    ctx.setup_powers(0x1234567, 70 * n, want_bytes=False)
    xy_host, _ = ctx.msm_g1(big, 70 * n, 0)
    ctx.msm_g1_launch(dev, 70 * n, 1, 0)
    xy_dev, _ = ctx.msm_fetch(1)
    assert xy_host == xy_dev
    ctx.setup_powers(0x1234567, n + 1, want_bytes=False)
    ctx.set_option("setup_g2", 1)
    for _ in range(2):
        out, infs = ctx.msm_g1_batch(host, n, batch)
        assert out == want and infs == want_inf
    assert infs[5] and not infs[4]
    for b in (0, 127, 128, 255, 256, 299):  # both sides of every pass boundary
        xy, inf = ctx.msm_g1(host[b * n * 32:(b + 1) * n * 32], n, 0)
        assert out[b * 96:(b + 1) * 96] == xy and not inf
