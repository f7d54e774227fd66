"""G1 MSMs through the direct-lookup table (csrc/msm_direct.cuh) against the bucket pipeline, 4097-point SRS.
    python tools/gpu_probe_direct.py [GiB ...]"""
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
budgets = [int(a) for a in sys.argv[1:]] or [40, 70]
for cname in ("bls12381", "bn254"):
    rng = np.random.default_rng(1)
    for gb in budgets:
        ctx = kzg_b200.Context(cname, 0)
        ctx.set_option("setup_g2", 0)
        ctx.set_option("profile", 1)
        ctx.set_option("direct_table_gb", gb)
        t0 = time.perf_counter()
        ctx.setup_powers(SECRET, 4097, want_bytes=False)
        key = "%s_%dgb" % (cname, gb)
        out[key + "_setup_ms"] = round((time.perf_counter() - t0) * 1e3, 1)
        sc = rng.integers(0, 256, size=(4096 * 4097, ctx.SB), dtype=np.uint8)
        sc[:, ctx.SB - 1] &= 0x0F
        raw = sc.tobytes()
        for direct in (gb, 0):
            ctx.set_option("direct_table_gb", direct)
            tag = key + ("_direct" if direct else "_buckets")
            for count in (2, 1025, 4097):
                r = ctx.msm_g1(raw[:count * ctx.SB], count)
                out["%s_count%d" % (tag, count)] = {"ms": round(best(lambda: ctx.msm_g1(raw[:count * ctx.SB], count)), 4),
                                                     "window_bits": ctx.msm_stats()["window_bits"], "phases": ctx.phase_times(), "xy": r[0][:8].hex()}
            for batch in (64, 4096):
                r = ctx.msm_g1_batch(raw[:batch * 4097 * ctx.SB], 4097, batch)
                ms = best(lambda: ctx.msm_g1_batch(raw[:batch * 4097 * ctx.SB], 4097, batch), 2)
                out["%s_batch%dx4097" % (tag, batch)] = {"ms": round(ms, 3), "commits_per_s": round(batch / ms * 1e3, 1), "xy": r[0][-8:].hex()}
        ctx.close()
print(json.dumps(out, indent=1))
