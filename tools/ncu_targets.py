"""Small workloads for ncu captures of the non-headline kernels (run under `ncu -k regex:<kernel> -c 1 --set full`).
    python tools/ncu_targets.py --what g2msm|setup|quotient|msm20|msm12 [--table-pad 1]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
sys.path.insert(0, ROOT)
import kzg_b200  # noqa: E402
from bench import SECRET, random_scalars  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", required=True)
    ap.add_argument("--table-pad", type=int, default=0)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    ctx = kzg_b200.Context("bls12381", 0)
    if a.what == "g2msm":      # msm_accumulate<Field2>: 2^16 G2 terms (wide enough to fill the machine)
        n = 1 << 16
        ctx.setup_powers(SECRET, n, want_bytes=False)
        raw = random_scalars(n, 3).view(np.uint8).tobytes()
        for _ in range(a.reps):
            ctx.msm_g2(raw, n)
    elif a.what == "setup":    # fixed_base_mul<Fp> / <Fp2>, 2^18 indices
        ctx.set_option("precompute", 0)
        for _ in range(a.reps):
            ctx.setup_powers(SECRET, 1 << 18, want_bytes=False)
    elif a.what == "quotient":  # poly_mul_trunc: n = 8192, k = 4096 (config 5)
        raw = random_scalars(8192, 5).view(np.uint8).tobytes()
        for _ in range(a.reps):
            ctx.poly_quotient(raw, 0, 4096)
    elif a.what.startswith("msm"):
        n = 1 << int(a.what[3:])
        ctx.set_option("setup_g2", 0)
        ctx.set_option("precompute", 1)
        ctx.set_option("table_pad", a.table_pad)
        ctx.setup_powers(SECRET, n, want_bytes=False)
        raw = random_scalars(n, 0x4B5A47).view(np.uint8).tobytes()
        for _ in range(a.reps):
            ctx.msm_g1(raw, n)
    ctx.close()


if __name__ == "__main__":
    main()
