"""Parity AT BASELINE SIZES against the reference itself (SURVEY.md 8(c) last paragraph, BASELINE.json configs 2 and 5).

tests/golden/<curve>/kat_large.json was produced by tools/gen_golden_large.py from the UNMODIFIED reference
(oracle/_ref/kzgref-<curve>): SRS of 5000 (BLS12-381) / 8200 (BN254) points derived from a secret in the fixture and
loaded by the reference, polynomials of degree 148 / 4096 / 4228 (testing/blob1.txt chunked as eth_blob_test does) /
8191, opened at k in {1, 4, 57, 128, 4096} points.  The CUDA path must regenerate that SRS byte for byte (sha256 of
the export_setup file), reproduce the coefficient bytes of the reference's polyfit, and then every
commit.serialize() / proof.serialize() / polyeval_G2(Z) byte string; where the prebuilt reference binary is present
its verify_proof must accept the GPU proofs and refute tampered data (the 3 refutations of the reference's tests).
Exact integer arithmetic: the bar is byte equality."""
import hashlib
import os
import subprocess
import sys
import tempfile

import pytest

from tests.golden_util import ROOT, chunk_values, load_large, ref_from_hex, stream_scalars

sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import kzg_b200  # noqa: E402

pytestmark = pytest.mark.gpu

BLOB_DIR = os.path.join(ROOT, "oracle", "_ref", "dropin")  # blob1.txt / blob2.txt travel with the prebuilt reference


def _poly_of(case, ctx):
    rec = case["recipe"]
    if rec["kind"] == "stream":
        return kzg_b200.Poly(ctx, stream_scalars(ctx.r, rec["seed"], rec["n"]))
    if rec["kind"] == "string":
        return kzg_b200.Poly.from_blob(ctx, [ord(ch) for ch in rec["text"]], rec["offset"])  # interpolation on the device
    path = os.path.join(BLOB_DIR, rec["file"])
    if not os.path.exists(path):
        return None
    data = ref_from_hex(open(path).read())
    assert hashlib.sha256(data).hexdigest() == rec["sha256_bytes"]
    ys = chunk_values(data, rec["chunk"])
    assert len(ys) == rec["chunks"]
    return kzg_b200.Poly.from_blob(ctx, ys, 0)


@pytest.mark.parametrize("cname", ["bls12381", "bn254"])
def test_reference_known_answers_at_baseline_sizes(cname):
    kat = load_large(cname)
    ctx = kzg_b200.Context(cname, 0)
    ts = kzg_b200.TrustedSetup.generate(cname, kat["srs_n"], int(kat["secret"], 16), ctx=ctx)
    srs_file = ts.export_setup()
    assert hashlib.sha256(srs_file).hexdigest() == kat["srs_sha256"], "device-generated SRS differs from the one the reference loaded"
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "kzgref-" + cname)
    have_ref = os.path.exists(ref_bin)
    checked = 0
    with tempfile.TemporaryDirectory() as td:
        f = lambda n: os.path.join(td, n)
        if have_ref:
            open(f("srs"), "wb").write(srs_file)
        for case in kat["cases"]:
            P = _poly_of(case, ctx)
            if P is None:
                continue
            assert len(P) == case["n_coeffs"], case["name"]
            assert hashlib.sha256(P.raw).hexdigest() == case["coeffs_sha256"], case["name"] + ": coefficients differ from the reference's polyfit"
            commit = ts.create_commit(P)
            assert commit.serialize().hex() == case["commit"], case["name"]
            for op in case["openings"]:
                off, k = op["off"], op["k"]
                proof = ts.create_proof(P, off, k)
                assert proof.serialize().hex() == op["proof"], (case["name"], off, k)
                ys = ts.evaluate(P, off, k)
                assert hashlib.sha256(ys).hexdigest() == op["ys_sha256"], (case["name"], off, k)
                if "g2msm_vanishing" in op:  # the verifier's G2 MSM over Z (config 5)
                    _, z = ctx.interpolate_range(ys, off)
                    assert hashlib.sha256(z).hexdigest() == op["vanishing_sha256"], (case["name"], off, k)
                    zxy, zinf = ctx.msm_g2(z, k + 1)
                    assert (b"\x04" + zxy).hex() == op["g2msm_vanishing"][8:], (case["name"], off, k)
                checked += 1
                # the reference's own verify_proof on the GPU's bytes (bounded: the large openings of the large cases)
                if have_ref and k in (57, 4096) and case["n_coeffs"] > 1000 and "verify" in op:
                    open(f("cm"), "wb").write(commit.serialize())
                    open(f("pf"), "wb").write(proof.serialize())
                    open(f("y"), "wb").write(ys)
                    run = lambda *a: subprocess.run([ref_bin, "verify", f("srs"), f("cm"), f("pf"), *map(str, a)], check=True,
                                                    capture_output=True, text=True).stdout.strip()
                    assert run(off, f("y")) == "1" == str(op["verify"])
                    t = bytearray(ys)
                    t[len(t) // 2] ^= 0x10
                    open(f("y2"), "wb").write(bytes(t))
                    assert run(off, f("y2")) == "0"          # tampered value
                    assert run(off + 1, f("y")) == "0"       # right values, wrong position
                    open(f("y3"), "wb").write(ys[:-ctx.SB])
                    assert run(off, f("y3")) == "0"          # one point short
    assert checked >= 14
    ctx.close()


def test_config5_nonzero_quotient_bn254():
    """BASELINE config 5 as SURVEY 8(d) states it: BN254, degree 8191, k = 4096, so the quotient has 4096 coefficients
    (the reference's own benchmark only ever opens a degree-4095 polynomial at 4096 points: q = 0)."""
    kat = load_large("bn254")
    case = [c for c in kat["cases"] if c["name"] == "deg8191_uniform"][0]
    op = [o for o in case["openings"] if o["k"] == 4096][0]
    ctx = kzg_b200.Context("bn254", 0)
    ts = kzg_b200.TrustedSetup.generate("bn254", kat["srs_n"], int(kat["secret"], 16), ctx=ctx)
    P = _poly_of(case, ctx)
    q = ctx.poly_quotient(P.raw, op["off"], 4096)
    assert len(q) == 4096 * ctx.SB and any(q)
    for mode in (3, 2, 1, 0):  # power series, cluster long division, one CTA, systolic: all the same quotient
        ctx.set_option("long_division", mode)
        assert ctx.poly_quotient(P.raw, op["off"], 4096) == q, mode
        assert ts.create_proof(P, op["off"], 4096).serialize().hex() == op["proof"], mode
    ctx.set_option("long_division", 3)
    xy, inf = ctx.msm_g1(q, 4096)   # proof = [q(s)]G1
    assert kzg_b200.Proof(xy, inf).serialize().hex() == op["proof"]
    ctx.close()
