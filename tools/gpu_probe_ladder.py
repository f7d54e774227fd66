"""Small MSMs against a LARGE resident SRS (2^20 points), with and without the table ladder (ctx.cuh TableRung).
    python tools/gpu_probe_ladder.py"""
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
ctx = kzg_b200.Context("bls12381", 0)
ctx.set_option("setup_g2", 0)
t0 = time.perf_counter()
ctx.setup_powers(SECRET, 1 << 20, want_bytes=False)
out["setup_2p20_ms"] = round((time.perf_counter() - t0) * 1e3, 2)
rng = np.random.default_rng(1)
sc = rng.integers(0, 256, size=(64 * 4097, ctx.SB), dtype=np.uint8)
sc[:, ctx.SB - 1] &= 0x0F
raw = sc.tobytes()
for ladder in (1, 0):
    ctx.set_option("table_ladder", ladder)
    for count in (2, 1025, 4097, 8192, 8193, 65536, 131072, 131073):
        r = ctx.msm_g1(raw[:count * ctx.SB], count)
        out["ladder%d_count%d" % (ladder, count)] = {"ms": round(best(lambda: ctx.msm_g1(raw[:count * ctx.SB], count)), 4),
                                                      "window_bits": ctx.msm_stats()["window_bits"], "xy": r[0][:8].hex()}
    r = ctx.msm_g1_batch(raw, 4097, 64)
    out["ladder%d_batch64x4097" % ladder] = {"ms": round(best(lambda: ctx.msm_g1_batch(raw, 4097, 64), 3), 4),
                                             "window_bits": ctx.msm_stats()["window_bits"], "xy": r[0][:8].hex()}
ctx.close()
print(json.dumps(out, indent=1))
