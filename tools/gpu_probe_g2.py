"""G2 (verify-side) MSM through both reduction tails, and the cost of making a 4096-point setup resident.
    python tools/gpu_probe_g2.py"""
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


def best(fn, reps=7):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t0) * 1e3)
    return min(ts)


out = {}
for cname in ("bls12381", "bn254"):
    ctx = kzg_b200.Context(cname, 0)
    ctx.set_option("profile", 1)
    ctx.setup_powers(SECRET, 5000, want_bytes=False)
    rng = np.random.default_rng(1)
    sc = rng.integers(0, 256, size=(4097, ctx.SB), dtype=np.uint8)
    sc[:, ctx.SB - 1] &= 0x0F
    raw = sc.tobytes()
    for proj in (2, 1):
        ctx.set_option("tail_proj", proj)
        for count in (2, 129, 1025, 4097):
            for grp, fn in (("g2", ctx.msm_g2), ("g1", ctx.msm_g1)):
                ms = best(lambda: fn(raw[:count * ctx.SB], count))
                out["%s_%s_%d_proj%d" % (cname, grp, count, proj)] = {"ms": round(ms, 4), "phases": ctx.phase_times()}
    ctx.set_option("tail_proj", 1)
    out[cname + "_latency_us"] = {name: round(1e6 / ctx.microbench(kind), 3) for kind, name in
                                  ((10, "g2_add_xyzz_16lane"), (11, "g2_add_xyzz_4lane"), (12, "g2_add_proj_32lane"),
                                   (13, "g1_add_proj_8lane"), (5, "g1_add_xyzz_4lane"), (7, "fp_modmul"))}
    ctx.close()
    ts = []
    for rep in range(4):
        c2 = kzg_b200.Context(cname, 0)
        t0 = time.perf_counter()
        c2.setup_powers(SECRET, 4096)
        ts.append((time.perf_counter() - t0) * 1e3)
        c2.close()
    out[cname + "_setup_4096_ms"] = ts
print(json.dumps(out, indent=1))
