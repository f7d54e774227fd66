import os, sys, time, random
sys.path.insert(0, "/root/repo/kzg-commitments_b200")
import kzg_b200
ctx = kzg_b200.Context("bls12381", 0)
rng = random.Random(1)
n = 4096
raw = b"".join(rng.randrange(ctx.r).to_bytes(ctx.SB, "little") for _ in range(n))
for k in (1, 2, 4, 8, 16, 32, 47, 48, 64, 96):
    ctx.poly_quotient(raw, 0, k)
    t0 = time.perf_counter()
    for _ in range(10):
        ctx.poly_quotient(raw, 0, k)
    print(f"k={k:4d} quotient {1e3 * (time.perf_counter() - t0) / 10:7.3f} ms", flush=True)
