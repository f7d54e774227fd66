"""Latency probes of the reduction tail (kzg_gpu_microbench kinds 4..7) and the effect of the barrier's poll pause on
small MSMs.   python tools/gpu_probe_latency.py > gpurun_out/latency.json"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
sys.path.insert(0, ROOT)
import kzg_b200  # noqa: E402
from bench import SECRET, random_scalars  # noqa: E402

ctx = kzg_b200.Context("bls12381", 0)
ctx.set_option("setup_g2", 0)
ctx.set_option("profile", 1)
stream = torch.cuda.ExternalStream(ctx.stream)
out = {"us_per_step": {}, "msm_ms": {}}
for sleep in (0, 20, 64, 200, 1000):
    ctx.set_option("barrier_sleep", sleep)
    out["us_per_step"]["grid_barrier_sleep%d" % sleep] = 1e6 / ctx.microbench(4)
for kind, name in ((5, "team_add"), (6, "thread_add"), (7, "fp_modmul"), (8, "fp_inv_divsteps"), (9, "fp_inv_bingcd")):
    out["us_per_step"][name] = 1e6 / ctx.microbench(kind)
for lg in (12, 16):
    n = 1 << lg
    ctx.setup_powers(SECRET, n, want_bytes=False)
    sc = torch.from_numpy(random_scalars(n, lg).view(np.uint8).reshape(-1)).cuda()
    for sleep in (0, 20, 64, 200):
        ctx.set_option("barrier_sleep", sleep)
        ts = []
        for rep in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            ctx.msm_g1_launch(sc, n, 1, 0)
            e1.record(stream)
            e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        out["msm_ms"]["2p%d_sleep%d" % (lg, sleep)] = {"ms": min(ts[1:]), "phases": ctx.phase_times()}
print(json.dumps(out, indent=1))
