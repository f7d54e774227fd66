"""Size-independent properties at BASELINE.json's FULL sizes, through the C ABI (include/kzg_gpu.h) with the CPU
oracle (oracle/kzg_oracle.py) as the checker only:

* config 4 — one BLS12-381 G1 MSM over 2^20 and 2^22 SRS points [s^i]G1 (polyeval_G1, reference
  src/trusted_setup.cpp:149-174): sum c_i [s^i]G = [P(s)]G with P(s) evaluated in Fr on the host, the SURVEY 8(d)
  edge set (all-zero, all-one, all r-1, single non-zero, 50 % zeros) at the same size, and the 8-way point-range
  split with the fold of the projective partials giving the bytes of the unsplit MSM;
* config 3 — 4096 blobs x 4096 coefficients (512 MiB of host scalars, kzg_gpu_msm_g1_batch, the call
  create_commit makes per blob, reference src/trusted_setup.cpp:137-142): a 64-blob subset against [P_b(s)]G and
  a checksum of checksums over ALL outputs, sum_b commit_b = [sum_b P_b(s)]G.

Exact integer arithmetic: the bar is byte equality."""
import os
import sys

import numpy as np
import pytest

from oracle.kzg_oracle import get_curve
from tests.golden_util import ROOT

sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import kzg_b200  # noqa: E402

pytestmark = pytest.mark.gpu

CNAME = "bls12381"
SECRET = 0x5EED0123456789ABCDEF0123456789ABCDEF0123456789ABCDEF


@pytest.fixture(scope="module")
def ctx():
    c = kzg_b200.Context(CNAME, 0)
    c.set_option("setup_g2", 0)  # G1 only: the G2 powers are not on this path
    yield c
    c.close()


def _random_records(n, seed):
    """n little-endian 32-byte scalars below 2^254 < r"""
    a = np.random.default_rng(seed).integers(0, 256, size=(n, 32), dtype=np.uint8)
    a[:, 31] &= 0x3F
    return a


def _ints(records):
    raw = records.tobytes()
    return [int.from_bytes(raw[i:i + 32], "little") for i in range(0, len(raw), 32)]


def _eval(C, coeffs, s):
    acc = 0
    for c in reversed(coeffs):
        acc = (acc * s + c) % C.r
    return acc


def _xy(C, k):
    return C.g1_to_octet(C.mul(k % C.r, C.g1) if k % C.r else None)[1:]


@pytest.mark.parametrize("log_n", [20, 22])
def test_single_msm_full_size(ctx, log_n):
    C = get_curve(CNAME)
    n = 1 << log_n
    s = SECRET % C.r
    ctx.setup_powers(s, n, want_bytes=False)
    rec = _random_records(n, 1000 + log_n)
    xy, inf = ctx.msm_g1(rec, n)
    assert not inf and xy == _xy(C, _eval(C, _ints(rec), s))

    # edge set of SURVEY 8(d) at the same size
    geo = (pow(s, n, C.r) - 1) * pow(s - 1, -1, C.r) % C.r  # sum of s^i, i < n
    ones = np.zeros((n, 32), dtype=np.uint8)
    ones[:, 0] = 1
    xy1, _ = ctx.msm_g1(ones, n)
    assert xy1 == _xy(C, geo)
    top = np.frombuffer((C.r - 1).to_bytes(32, "little"), dtype=np.uint8)
    xym, _ = ctx.msm_g1(np.ascontiguousarray(np.broadcast_to(top, (n, 32))), n)
    assert xym == _xy(C, -geo)
    xy0, inf0 = ctx.msm_g1(np.zeros((n, 32), dtype=np.uint8), n)
    assert inf0 and xy0 == bytes(96)
    single = np.zeros((n, 32), dtype=np.uint8)
    single[n - 1] = rec[7]
    xys, _ = ctx.msm_g1(single, n)
    assert xys == _xy(C, int.from_bytes(rec[7].tobytes(), "little") * pow(s, n - 1, C.r))
    half = rec.copy()
    half[np.random.default_rng(5).integers(0, 2, size=n).astype(bool)] = 0
    xyh, _ = ctx.msm_g1(half, n)
    assert xyh == _xy(C, _eval(C, _ints(half), s))

    # config 4: contiguous point-range split over 8 shards, projective partials folded as after the all-gather
    ctx.set_option("finish_affine", 0)
    partials = []
    for g in range(8):
        lo, hi = g * n // 8, (g + 1) * n // 8
        ctx.msm_g1(np.ascontiguousarray(rec[lo:hi]), hi - lo, base_offset=lo)
        partials.append(ctx.msm_fetch_partial())
    ctx.set_option("finish_affine", 1)
    xyf, inff = ctx.g1_sum_partials(partials)
    assert not inff and xyf == xy


def test_blob_batch_4096_x_4096(ctx):
    C = get_curve(CNAME)
    n = batch = 4096
    s = SECRET % C.r
    ctx.setup_powers(s, n + 1, want_bytes=False)
    rec = _random_records(batch * n, 3)
    rec[9 * n:10 * n] = 0  # one all-zero blob -> infinity
    out, infs = ctx.msm_g1_batch(rec, n, batch)
    assert infs[9] and sum(infs) == 1

    # 64-blob subset against [P_b(s)]G
    evals = {}
    for b in list(range(0, batch, 65)) + [batch - 1]:
        evals[b] = _eval(C, _ints(rec[b * n:(b + 1) * n]), s)
        assert out[b * 96:(b + 1) * 96] == _xy(C, evals[b]), b

    # checksum of checksums over all outputs: sum_b commit_b = [ (sum_b P_b)(s) ]G, the coefficient sums taken per
    # 32-bit limb column in uint64 (4096 x 2^32 < 2^44)
    limbs = rec.reshape(batch, n, 8, 4).view("<u4").reshape(batch, n, 8).sum(axis=0, dtype=np.uint64)
    summed = [sum(int(limbs[i, j]) << (32 * j) for j in range(8)) % C.r for i in range(n)]
    total = None
    for b in range(batch):
        if infs[b]:
            continue
        P = (int.from_bytes(out[b * 96:b * 96 + 48], "big"), int.from_bytes(out[b * 96 + 48:(b + 1) * 96], "big"))
        total = C.add(total, P)
    assert C.g1_to_octet(total)[1:] == _xy(C, _eval(C, summed, s))
    # and the summed polynomial committed on the device
    xy, _ = ctx.msm_g1(C.scalars_to_bytes(summed), n)
    assert xy == C.g1_to_octet(total)[1:]


def test_concurrent_callers_on_one_context(ctx):
    """SURVEY 8(b) threading: create_commit / create_proof only read the SRS and may be called from several threads
    (reference src/kzg.h:182-290 has no locks); the context serialises the shared scratch internally.  Four threads
    issue commitments, proofs and quotients of different sizes on ONE context at once (ctypes drops the GIL for the
    duration of each call): every result must equal the one computed alone."""
    import threading
    C = get_curve(CNAME)
    s = SECRET % C.r
    ctx.set_option("setup_g2", 0)
    ctx.setup_powers(s, 5000, want_bytes=False)
    jobs = []
    for t in range(4):
        n = (4097, 1500, 300, 4999)[t]
        rec = _random_records(n, 50 + t)
        jobs.append((n, rec, rec.tobytes()))
    alone = [(ctx.msm_g1(rec, n), ctx.prove(raw, 3 + i, 2), ctx.poly_quotient(raw, 1, 5))
             for i, (n, rec, raw) in enumerate(jobs)]
    assert alone[0][0][0] == _xy(C, _eval(C, _ints(jobs[0][1]), s))
    errors = []

    def worker(i):
        n, rec, raw = jobs[i]
        try:
            for _ in range(25):
                got = (ctx.msm_g1(rec, n), ctx.prove(raw, 3 + i, 2), ctx.poly_quotient(raw, 1, 5))
                if got != alone[i]:
                    errors.append((i, "mismatch"))
                    return
        except Exception as e:  # noqa: BLE001
            errors.append((i, repr(e)))

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(4)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors
