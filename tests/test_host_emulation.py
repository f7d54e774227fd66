"""CPU tests of the CUDA sources themselves, compiled for the host (kzg-commitments_b200/csrc/host_check.cpp):
the PTX carry chains run through the flag emulation of csrc/ptx.cuh and the per-thread kernel bodies of
csrc/msm.cuh are driven by loops.  Everything is compared with the oracle / the reference's golden vectors.
These tests exercise the same C++ the GPU runs, but they are NOT the parity tests proper (tests/test_gpu_*)."""
import ctypes
import os
import random
import subprocess

import pytest

from oracle.kzg_oracle import get_curve
from tests.golden_util import CURVE_NAMES, load

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "kzg-commitments_b200")


@pytest.fixture(scope="module")
def hc():
    so = os.path.join(PKG, "libkzg_hostcheck.so")
    subprocess.run(["make", "-C", PKG, "libkzg_hostcheck.so"], check=True, capture_output=True)
    lib = ctypes.CDLL(so)
    lib.hc_msm.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_uint64, ctypes.c_char_p, ctypes.c_char_p,
                           ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, ctypes.c_int,
                           ctypes.c_int, ctypes.c_int, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int)]
    lib.hc_selftest.argtypes = [ctypes.c_int, ctypes.c_uint32]
    lib.hc_msm_direct.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_uint64,
                                  ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_char_p,
                                  ctypes.POINTER(ctypes.c_int)]
    return lib


def _words(v, n):
    return (ctypes.c_uint32 * n)(*[(v >> (32 * i)) & 0xFFFFFFFF for i in range(n)])


def _val(arr):
    return sum(int(w) << (32 * i) for i, w in enumerate(arr))


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_selftest_identities(hc, cname):
    C = get_curve(cname)
    for seed in range(3):
        assert hc.hc_selftest(C.id, seed) == 0


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_montgomery_field_ops(hc, cname):
    C = get_curve(cname)
    rng = random.Random(7 + C.id)
    for which, mod in ((0, C.p), (1, C.r)):
        n = hc.hc_limbs(C.id, which)
        edge = [0, 1, 2, mod - 1, mod - 2, (1 << 32) - 1, 1 << 32, (1 << (mod.bit_length() - 1))]
        vals = edge + [rng.randrange(mod) for _ in range(40)]
        for a in vals:
            b = rng.choice(vals)
            out = (ctypes.c_uint32 * n)()
            for op, want in ((0, a * b % mod), (1, (a + b) % mod), (2, (a - b) % mod), (4, -a % mod)):
                assert hc.hc_field_op(C.id, which, op, _words(a, n), _words(b, n), out) == 0
                assert _val(out) == want, (cname, which, op, hex(a), hex(b))
        for a in vals:
            out = (ctypes.c_uint32 * n)()
            hc.hc_field_op(C.id, which, 3, _words(a, n), _words(0, n), out)
            assert _val(out) == (pow(a, -1, mod) if a else 0)
        # non-canonical inputs (full limb range) are reduced by to_mont
        a = (1 << (32 * n)) - 1
        out = (ctypes.c_uint32 * n)()
        hc.hc_field_op(C.id, which, 0, _words(a, n), _words(1, n), out)
        assert _val(out) == a % mod


def _pts_bytes(C, pts, g2):
    return b"".join((C.g2_to_octet(P) if g2 else C.g1_to_octet(P))[1:] for P in pts)


def _run(hc, C, group, pts, scalars, count, base_offset=0, batch=1, c=4, chunk=3, bpc=2, precomp=0):
    g2 = group == 2
    nco = 4 if g2 else 2
    out = ctypes.create_string_buffer(batch * nco * C.M)
    inf = (ctypes.c_int * batch)()
    rc = hc.hc_msm(C.id, group, len(pts), _pts_bytes(C, pts, g2), scalars, count, batch, base_offset, c, chunk, bpc,
                   precomp, out, inf)
    assert rc == 0
    return out.raw, list(inf)


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_msm_pipeline_matches_reference_commits(hc, cname):
    """The host-run kernel bodies reproduce commit.serialize() / polyeval_G2 of the REFERENCE (golden vectors)."""
    C = get_curve(cname)
    kat, srs = load(cname)
    g1, g2 = C.srs_from_file_bytes(srs)
    for case in kat["cases"]:
        coeffs = bytes.fromhex(case["coeffs"])
        n = len(coeffs) // C.SB
        if n == 0:
            continue
        for c, chunk, bpc, precomp in ((4, 3, 2, 0), (5, 64, 0, 0), (6, 2, 1, 1)):  # bpc = affine pre-reduction levels
            raw, inf = _run(hc, C, 1, g1, coeffs, n, c=c, chunk=chunk, bpc=bpc, precomp=precomp)
            want = bytes.fromhex(case["commit"])[5:]
            assert raw == want, (cname, case["name"], c, precomp)
            assert inf[0] == int(want == bytes(2 * C.M))
        raw, inf = _run(hc, C, 2, g2, coeffs, n, c=4)
        assert raw == bytes.fromhex(case["g2msm"])[5:], (cname, case["name"], "g2")


@pytest.mark.parametrize("cname", ["bls12381", "bn158"])
def test_msm_pipeline_edges(hc, cname):
    C = get_curve(cname)
    rng = random.Random(99)
    # repeated bases force P+P (doubling) and P-P (infinity) inside bucket accumulation and reduction
    P = C.mul(5, C.g1)
    pts = [P] * 6 + [C.neg(P)] * 2 + [None, C.g1]
    cases = [
        [1] * 10,
        [3, 3, 3, 3, 3, 3, 3, 3, 0, 0],
        [C.r - 1] * 10,
        [7, 7, C.r - 7, C.r - 7, 0, 0, 7, 7, 5, 0],     # cancels to infinity except the trailing terms
        [rng.randrange(C.r) for _ in range(10)],
        [0] * 10,
        [2 ** 255 % (1 << (8 * C.SB)), (1 << (8 * C.SB)) - 1] + [0] * 8,   # non-canonical scalars are reduced mod r
    ]
    for sc in cases:
        raw_sc = b"".join((s % (1 << (8 * C.SB))).to_bytes(C.SB, "little") for s in sc)
        want = C.msm([s % C.r for s in sc], pts)
        for c, precomp in ((4, 0), (7, 0), (5, 1)):
            raw, inf = _run(hc, C, 1, pts, raw_sc, len(sc), c=c, chunk=2, bpc=3, precomp=precomp)
            assert raw == C.g1_to_octet(want)[1:], (sc, c, precomp)
            assert inf[0] == int(want is None)
    # base_offset and batch: two MSMs over a sub-range
    kat, srs = load(cname)
    g1, _ = C.srs_from_file_bytes(srs)
    sc = [rng.randrange(C.r) for _ in range(2 * 9)]
    raw, inf = _run(hc, C, 1, g1, C.scalars_to_bytes(sc), 9, base_offset=0, batch=2, c=5, chunk=4, bpc=1)
    for b in range(2):
        want = C.msm(sc[9 * b:9 * b + 9], g1[:9])
        assert raw[b * 2 * C.M:(b + 1) * 2 * C.M] == C.g1_to_octet(want)[1:]
    raw, inf = _run(hc, C, 1, g1, C.scalars_to_bytes(sc[:9]), 9, base_offset=11, c=5, precomp=1)
    assert raw == C.g1_to_octet(C.msm(sc[:9], g1[11:20]))[1:]


@pytest.mark.parametrize("cname", ["bls12381", "bn254"])
def test_direct_lookup_path(hc, cname):
    """csrc/msm_direct.cuh on the host: the table of every digit multiple, one partial sum per (scalars, window group)
    slot (group-major, padded to warps), pairwise trees — against the oracle's MSM; repeated and opposite bases, an
    infinity in the SRS, point ranges, batches, non-canonical scalars."""
    C = get_curve(cname)
    rng = random.Random(2024 + C.id)
    P = C.mul(11, C.g1)
    pts = [C.g1, P, P, C.neg(P), None, C.mul(rng.randrange(C.r), C.g1), C.mul(7, C.g1)]
    raw_pts = _pts_bytes(C, pts, False)

    def run(scalars, count, batch=1, off=0, c=3, wpt=4, spt=1):
        raw_sc = b"".join((s % (1 << (8 * C.SB))).to_bytes(C.SB, "little") for s in scalars)
        out = ctypes.create_string_buffer(batch * 2 * C.M)
        inf = (ctypes.c_int * batch)()
        assert hc.hc_msm_direct(C.id, len(pts), raw_pts, raw_sc, count, batch, off, c, wpt, spt, out, inf) == 0
        return out.raw, list(inf)

    n = len(pts)
    for sc in ([rng.randrange(C.r) for _ in range(n)], [1] * n, [C.r - 1] * n, [0] * n,
               [5, 3, C.r - 3, 0, 9, 0, 0], [(1 << (8 * C.SB)) - 1] + [0] * (n - 1)):
        want = C.msm([s % C.r for s in sc], pts)
        for c, wpt, spt in ((3, 4, 1), (4, 1000, 2), (2, 7, 1), (5, 1000, 1)):   # several window groups / all windows in one slot
            raw, inf = run(sc, n, c=c, wpt=wpt, spt=spt)
            assert raw == C.g1_to_octet(want)[1:] and inf[0] == int(want is None), (sc[:2], c, wpt, spt)
    # a point range and a batch of two
    sc = [rng.randrange(C.r) for _ in range(6)]
    raw, inf = run(sc, 3, batch=2, off=2, c=3, wpt=5)
    for b in range(2):
        want = C.msm(sc[3 * b:3 * b + 3], pts[2:5])
        assert raw[b * 2 * C.M:(b + 1) * 2 * C.M] == C.g1_to_octet(want)[1:] and inf[b] == int(want is None), b
