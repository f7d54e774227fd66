"""GPU probe (development tool): wall time of kzg_gpu_interpolate_range (poly::from_blob / verify's I and Z)."""
import os, sys, time, random
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import kzg_b200
for curve in ("bls12381", "bn254"):
    ctx = kzg_b200.Context(curve, 0)
    rng = random.Random(1)
    for k in (128, 512, 1024, 2048, 4096, 8192):
        ys = b"".join(rng.randrange(ctx.r).to_bytes(ctx.SB, "little") for _ in range(k))
        ctx.interpolate_range(ys, 0)
        t0 = time.perf_counter()
        for _ in range(5):
            ctx.interpolate_range(ys, 0)
        t_both = (time.perf_counter() - t0) / 5 * 1e3
        t0 = time.perf_counter()
        for _ in range(5):
            ctx.interpolate_range(ys, 0, want_vanishing=False)
        t_i = (time.perf_counter() - t0) / 5 * 1e3
        print(f"{curve} k={k:5d} interpolate I+Z {t_both:7.3f} ms   I only {t_i:7.3f} ms", flush=True)
    ctx.close()
