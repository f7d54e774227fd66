"""kzg_gpu_poly_eval_range_noalloc (include/kzg_gpu.h): the entry point that allocates nothing — scratch in a static
__device__ array, device buffers in and out — against the existing kzg_gpu_poly_eval_range, through the C ABI.
Exact integer arithmetic: the tolerance is byte equality, as everywhere else in this suite.  Two calls in a row (the
second with smaller sizes, so it runs on top of the first call's leftovers in the shared array), the size limit, and a
second context on the same device (same array).  Collected last on purpose (file name)."""
import os
import random
import sys

import pytest

from oracle.kzg_oracle import get_curve
from tests.golden_util import CURVE_NAMES, ROOT

sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import kzg_b200  # noqa: E402

pytestmark = pytest.mark.gpu

MAX_COEFFS = 65536  # KZG_GPU_NOALLOC_MAX_COEFFS
MAX_POINTS = 65536  # KZG_GPU_NOALLOC_MAX_POINTS


def _noalloc_eval(ctx, raw, first_root, k):
    import torch
    n = len(raw) // ctx.SB
    d_in = torch.frombuffer(bytearray(raw), dtype=torch.uint8).cuda()
    d_out = torch.full((k * ctx.SB,), 0xAB, dtype=torch.uint8, device="cuda")  # the caller's buffers, made before the call
    torch.cuda.synchronize()
    ctx.poly_eval_range_noalloc(d_in, n, first_root, k, d_out)
    torch.cuda.ExternalStream(ctx.stream).synchronize()  # the call only enqueues on the context's stream
    return bytes(d_out.cpu().numpy().tobytes())


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_noalloc_eval_matches_existing_entry_point_twice_in_a_row(cname):
    C = get_curve(cname)
    ctx = kzg_b200.Context(cname, 0)
    rng = random.Random(0x20A110C + C.id)
    for n, first_root, k in ((5000, 3, 700), (300, -2, 40), (1, 0, 1)):  # back to back; later calls are smaller
        coeffs = [rng.randrange(C.r) for _ in range(n)]
        raw = C.scalars_to_bytes(coeffs)
        want = ctx.poly_eval_range(raw, first_root, k)          # the existing entry point (host buffers, its own scratch)
        assert _noalloc_eval(ctx, raw, first_root, k) == want, (cname, n, k)
    assert want == C.scalars_to_bytes([coeffs[0]])              # and the oracle, for good measure
    # a second context on the same device shares the same static array
    other = kzg_b200.Context(cname, 0)
    raw = C.scalars_to_bytes([rng.randrange(C.r) for _ in range(1234)])
    assert _noalloc_eval(other, raw, 10, 77) == ctx.poly_eval_range(raw, 10, 77)
    assert _noalloc_eval(ctx, raw, 10, 77) == ctx.poly_eval_range(raw, 10, 77)
    other.close()
    ctx.close()


def test_noalloc_limits_are_rejected_with_a_message():
    import torch
    ctx = kzg_b200.Context("bls12381", 0)
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    for n, k in ((MAX_COEFFS + 1, 1), (1, MAX_POINTS + 1)):
        with pytest.raises(RuntimeError, match="exceed the static scratch"):
            ctx.poly_eval_range_noalloc(d, n, 0, k, d)
    with pytest.raises(RuntimeError):
        ctx.poly_eval_range_noalloc(0, 4, 0, 1, d)  # null input
    # at the limit itself it runs: 65536 coefficients, a few points
    C = get_curve("bls12381")
    rng = random.Random(9)
    raw = C.scalars_to_bytes([rng.randrange(C.r) for _ in range(MAX_COEFFS)])
    assert _noalloc_eval(ctx, raw, 1, 5) == ctx.poly_eval_range(raw, 1, 5)
    ctx.close()
