"""GPU probe (development tool, run under gpurun): integer-pipe microbenchmarks and a sweep of the G1 MSM
over sizes / window bits / task chunk, timed with CUDA events on the context stream.
    python tools/gpu_probe.py [--sizes 16,18,20] [--windows 0,13,14,15,16] [--out gpurun_out/probe.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import kzg_b200  # noqa: E402


def random_scalars(n, seed, r=0x73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001, sb=32):
    """n uniform scalars below (top 64 bits of r) << (8*sb - 64), i.e. below r, as an (n, sb) uint8 array (little-endian)."""
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, size=(n, sb), dtype=np.uint8)
    top = rng.integers(0, r >> (8 * (sb - 8)), size=n, dtype=np.uint64)
    a[:, sb - 8:] = top.view(np.uint8).reshape(n, 8)
    return a


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--curve", default="bls12381")
    ap.add_argument("--sizes", default="16,18,20")
    ap.add_argument("--windows", default="0")
    ap.add_argument("--chunks", default="0")
    ap.add_argument("--precompute", default="0")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--team", default="1")
    ap.add_argument("--table-pad", type=int, default=0)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    ctx = kzg_b200.Context(args.curve, 0)
    res = {"microbench": {}, "msm": []}
    for kind, name in ((0, "wide_mad_products_per_s"), (1, "imad32_per_s"), (2, "fp_modmul_per_s"), (3, "fr_modmul_per_s")):
        res["microbench"][name] = ctx.microbench(kind)
    print(json.dumps(res["microbench"]), flush=True)
    stream = torch.cuda.ExternalStream(ctx.stream)
    ctx.set_option("setup_g2", 0)
    ctx.set_option("profile", 1)
    ctx.set_option("table_pad", args.table_pad)
    for lg in [int(x) for x in args.sizes.split(",")]:
        n = 1 << lg
        sc = torch.from_numpy(random_scalars(n, lg, ctx.r, ctx.SB).reshape(-1)).cuda()
        for pre in [int(x) for x in args.precompute.split(",")]:
            for wb in [int(x) for x in args.windows.split(",")]:
                ctx.set_option("precompute", pre)
                ctx.set_option("window_bits", wb)
                t0 = time.time()
                ctx.setup_powers(12345678901234567890123, n, want_bytes=False)
                t_setup = time.time() - t0
                for ch, team in [(int(x), int(y)) for x in args.chunks.split(",") for y in args.team.split(",")]:
                    ctx.set_option("task_chunk", ch)
                    ctx.set_option("team", team)
                    times = []
                    for rep in range(args.reps + 1):
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record(stream)
                        ctx.msm_g1_launch(sc, n, 1, 0)
                        e1.record(stream)
                        e1.synchronize()
                        times.append(e0.elapsed_time(e1))
                    xy, inf = ctx.msm_fetch(1)
                    row = {"log2n": lg, "precompute": pre, "window_bits": wb, "task_chunk": ch, "team": team, "ms": min(times[1:]),
                           "first_ms": times[0], "mpts_per_s": n / min(times[1:]) / 1e3, "setup_s": t_setup,
                           "xy": xy[:8].hex(), "phases": ctx.phase_times(), "stats": ctx.msm_stats()}
                    res["msm"].append(row)
                    print(json.dumps(row), flush=True)
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
