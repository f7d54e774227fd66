"""create_proof on LARGE-DEGREE polynomials (the reference's --benchmark-common shape, benchmark/benchmark.cpp:123-136):
kzg_gpu_prove timed through the C ABI (host coefficients in, proof bytes out) and checked against the oracle through
the known secret: proof = [q(s)]G1 with q(s) = (P(s) - I(s)) / Z(s); for k = 1, I = P(a) and Z = s - a; for larger k
the check evaluates q itself (returned by kzg_gpu_poly_quotient) at s in Fr.
    python tools/gpu_probe_large_degree.py > gpurun_out/large_degree.json"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
sys.path.insert(0, ROOT)
import kzg_b200  # noqa: E402
from bench import SECRET, fr_eval, random_scalars, scalars_to_ints  # noqa: E402
from oracle.kzg_oracle import get_curve  # noqa: E402  (checker only)

C = get_curve("bls12381")
s = SECRET % C.r
ctx = kzg_b200.Context("bls12381", 0)
ctx.set_option("setup_g2", 0)
out = []
for n, k, off in ((1 << 20, 1, 12345), (1 << 20, 64, 7), (65536, 32768, 0), (1 << 22, 1, 3)):
    ctx.setup_powers(SECRET, n, want_bytes=False)
    arr = random_scalars(n, n + k)
    raw = arr.view(np.uint8).tobytes()
    ctx.prove(raw, off, k)
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        xy, inf = ctx.prove(raw, off, k)
        ts.append((time.perf_counter() - t0) * 1e3)
    t0 = time.perf_counter()
    q = ctx.poly_quotient(raw, off, k)
    t_q = (time.perf_counter() - t0) * 1e3
    qs = fr_eval([int.from_bytes(q[i:i + 32], "little") for i in range(0, len(q), 32)], s, C.r)
    ok_q = xy == C.g1_to_octet(C.mul(qs, C.g1))[1:]
    ok_direct = None
    if k == 1:  # independent of our quotient kernel
        ints = scalars_to_ints(arr)
        ps, pa = fr_eval(ints, s, C.r), fr_eval(ints, off, C.r)
        ok_direct = xy == C.g1_to_octet(C.mul((ps - pa) * pow(s - off, -1, C.r) % C.r, C.g1))[1:]
    row = {"n": n, "k": k, "first_root": off, "prove_ms": min(ts), "quotient_call_ms": t_q, "proof_equals_[q(s)]G1": ok_q,
           "proof_equals_[(P(s)-P(a))/(s-a)]G1": ok_direct}
    out.append(row)
    print(json.dumps(row), file=sys.stderr, flush=True)
    assert ok_q and ok_direct is not False
print(json.dumps(out, indent=1))
