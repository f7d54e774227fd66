"""Development probe: G2 MSM (verify_proof's polyeval_G2) and setup timings."""
import os, sys, time, random
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import kzg_b200
for curve in ("bls12381", "bn254"):
    ctx = kzg_b200.Context(curve, 0)
    ctx.set_option("profile", 1)
    rng = random.Random(1)
    n = 4200
    for pre in (1, 2):
        ctx.set_option("precompute", pre)
        t0 = time.time(); ctx.setup_powers(rng.randrange(ctx.r), n, want_bytes=False); t_setup = time.time() - t0
        sc = b"".join(rng.randrange(ctx.r).to_bytes(ctx.SB, "little") for _ in range(4097))
        for rep in range(3):
            t0 = time.time(); ctx.msm_g2(sc, 4097); t = time.time() - t0
            print(curve, "precompute", pre, "g2 msm 4097: %.3f ms" % (t * 1e3), ctx.phase_times(), ctx.msm_stats(), "setup %.1f ms" % (t_setup * 1e3), flush=True)
        t0 = time.time(); ctx.msm_g1(sc, 4097); t = time.time() - t0
        print(curve, "g1 msm 4097: %.3f ms" % (t * 1e3), ctx.phase_times(), flush=True)
    ctx.close()
