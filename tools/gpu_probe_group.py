"""One process driving G devices through the kzg_gpu_group C ABI (include/kzg_gpu.h "Multi-GPU"): batched blob
commitments sharded by blob and one large MSM split by point range, timed on the host clock around the group call
and compared with the same call on a one-device group.
    python tools/gpu_probe_group.py --devices 2"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
sys.path.insert(0, ROOT)
import kzg_b200  # noqa: E402
from bench import SECRET, random_scalars  # noqa: E402


def best(fn, reps=3):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        r = fn()
        ts.append((time.perf_counter() - t0) * 1e3)
    return min(ts), r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--devices", type=int, default=2)
    ap.add_argument("--blobs", type=int, default=4096)
    args = ap.parse_args()
    import torch
    out = {"devices": args.devices}
    nc = 4096
    blobs = torch.from_numpy(random_scalars(args.blobs * nc, 7).view(np.uint8).reshape(-1)).pin_memory()
    big = torch.from_numpy(random_scalars(1 << 20, 9).view(np.uint8).reshape(-1)).pin_memory()
    res = {}
    for g in (1, args.devices):
        grp = kzg_b200.Group("bls12381", g)
        grp.set_option("setup_g2", 0)
        grp.setup_powers(SECRET, nc + 1, want_bytes=False)
        ms, r = best(lambda: grp.msm_g1_batch(blobs, nc, args.blobs))
        res[("batch", g)] = (ms, r[0])
        roots = [i % 4096 for i in range(args.blobs)]
        ms, r = best(lambda: grp.prove_batch(blobs, nc, roots, 1), reps=2)
        res[("prove", g)] = (ms, r[0])
        grp.setup_powers(SECRET, 1 << 20, want_bytes=False)
        ms, r = best(lambda: grp.msm_g1(big, 1 << 20))
        res[("msm", g)] = (ms, r[0])
        grp.close()
    G = args.devices
    for what in ("batch", "prove", "msm"):
        assert res[(what, 1)][1] == res[(what, G)][1], what + ": multi-device bytes differ from single-device bytes"
        out[what] = {"ms_1": res[(what, 1)][0], "ms_%d" % G: res[(what, G)][0], "speedup": res[(what, 1)][0] / res[(what, G)][0],
                     "bytes_identical": True}
    out["batch"]["commits_per_s"] = args.blobs / (res[("batch", G)][0] * 1e-3)
    out["prove"]["proofs_per_s"] = args.blobs / (res[("prove", G)][0] * 1e-3)
    out["msm"]["points_per_s"] = (1 << 20) / (res[("msm", G)][0] * 1e-3)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
