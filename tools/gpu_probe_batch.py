import os, sys, time
import numpy as np, torch
sys.path.insert(0, "/root/repo/kzg-commitments_b200"); sys.path.insert(0, "/root/repo/tools")
import kzg_b200
from gpu_probe import random_scalars
ctx = kzg_b200.Context("bls12381", 0)
ctx.set_option("setup_g2", 0); ctx.set_option("profile", 1); ctx.set_option("precompute", 1)
ctx.setup_powers(12345678901234567890123, 4097, want_bytes=False)
stream = torch.cuda.ExternalStream(ctx.stream)
for nb in (64, 256, 1024):
    sc = torch.from_numpy(random_scalars(nb * 4096, nb, ctx.r, ctx.SB).reshape(-1)).cuda()
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); ctx.msm_g1_launch(sc, 4096, nb, 0); e1.record(stream); e1.synchronize()
    print(nb, "blobs:", round(e0.elapsed_time(e1), 3), "ms", {k: round(v, 3) for k, v in ctx.phase_times().items()}, ctx.msm_stats(), flush=True)
