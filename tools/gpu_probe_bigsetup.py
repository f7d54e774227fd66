"""commit / prove(k = 1) of growing degree against one large resident SRS (what benchmark.cpp --benchmark-common does),
first and repeated calls.    python tools/gpu_probe_bigsetup.py"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
sys.path.insert(0, ROOT)
import kzg_b200  # noqa: E402
from bench import SECRET  # noqa: E402

out = {}
ctx = kzg_b200.Context("bls12381", 0)
t0 = time.perf_counter()
ctx.setup_powers(SECRET, 5 << 20, want_bytes=False)
out["setup_5M_ms"] = round((time.perf_counter() - t0) * 1e3, 1)
rng = np.random.default_rng(1)
for lg in (12, 16, 19, 20):
    n = (1 << lg) + 1
    sc = rng.integers(0, 256, size=(n, ctx.SB), dtype=np.uint8)
    sc[:, ctx.SB - 1] &= 0x0F
    raw = sc.tobytes()
    row = {}
    for what, fn in (("commit", lambda: ctx.msm_g1(raw, n)), ("prove", lambda: ctx.prove(raw, 0, 1))):
        ts = []
        for _ in range(3):
            t0 = time.perf_counter()
            fn()
            ts.append(round((time.perf_counter() - t0) * 1e3, 3))
        row[what] = ts
    out["degree_2p%d" % lg] = row
ctx.close()
print(json.dumps(out, indent=1))
