"""GPU probe (development tool): wall time of create_proof's device chain (kzg_gpu_prove) and of the quotient alone
for the multi-proof shapes of benchmark/benchmark.cpp (degree 4095, k = 128 .. 2048), per division kernel."""
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import kzg_b200  # noqa: E402

ctx = kzg_b200.Context("bls12381", 0)
ctx.setup_powers(123456789, 5000, want_bytes=False)
rng = random.Random(1)
n = 4096
raw = b"".join(rng.randrange(ctx.r).to_bytes(ctx.SB, "little") for _ in range(n))
for mode in (3, 2):
    ctx.set_option("long_division", mode)
    for k in (128, 256, 512, 1024, 2048):
        ctx.prove(raw, 0, k)
        t0 = time.perf_counter()
        for _ in range(5):
            ctx.prove(raw, 0, k)
        t_prove = (time.perf_counter() - t0) / 5 * 1e3
        t0 = time.perf_counter()
        for _ in range(5):
            ctx.poly_quotient(raw, 0, k)
        t_q = (time.perf_counter() - t0) / 5 * 1e3
        print(f"long_division={mode} k={k:5d} prove {t_prove:7.3f} ms   quotient (H2D + division + D2H) {t_q:7.3f} ms", flush=True)

ctx.set_option("long_division", 2)
for nn in (1025, 2049, 4097):
    sc = raw[:32] * nn
    ctx.msm_g1(sc, nn, 0)
    t0 = time.perf_counter()
    for _ in range(10):
        ctx.msm_g1(sc, nn, 0)
    print(f"msm_g1 host scalars n={nn}: {(time.perf_counter() - t0) / 10 * 1e3:.3f} ms  stats {ctx.msm_stats()}", flush=True)
