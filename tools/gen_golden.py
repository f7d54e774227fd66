#!/usr/bin/env python3
"""Generate tests/golden/<curve>/ known-answer vectors by running the UNMODIFIED reference
(oracle/_ref/kzgref-<curve>, built from /root/reference by oracle/build_ref.sh).

The reference's own tests hold no stored vectors (every test draws a fresh random SRS),
so these files are what pins parity: one SRS exported by the reference's `trusted_setup(int)` +
`export_setup`, and for a list of polynomials / openings the bytes of commit.serialize(),
proof.serialize(), polyeval_G2, the quotient, evaluations, Z, I and poly.serialize().

Run (in the container that has /root/reference):  python tools/gen_golden.py
"""
import json
import os
import random
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.kzg_oracle import get_curve  # noqa: E402  (only for r and byte packing)

SRS_N = 48


def run(binary, *args):
    return subprocess.run([binary, *map(str, args)], check=True, capture_output=True, text=True).stdout


def main():
    for cname in ("bls12381", "bn254", "bn158"):
        binary = os.path.join(ROOT, "oracle", "_ref", f"kzgref-{cname}")
        C = get_curve(cname)
        rng = random.Random(0x4B5A47 + C.id)
        outdir = os.path.join(ROOT, "tests", "golden", cname)
        os.makedirs(outdir, exist_ok=True)
        srs = os.path.join(outdir, "srs.bin")
        if not os.path.exists(srs):
            run(binary, "setup", SRS_N, srs)
        polys = {
            "zero": [],
            "const": [rng.randrange(C.r)],
            "deg1": [rng.randrange(C.r) for _ in range(2)],
            "deg10": [rng.randrange(C.r) for _ in range(11)],
            "deg37": [rng.randrange(C.r) for _ in range(38)],
            "deg46_max": [rng.randrange(C.r) for _ in range(SRS_N - 1)],
            "edge_scalars": [0, 1, C.r - 1, 2, C.r - 2, 0, 0, 1, 1, (1 << 128) - 1, 1 << 128, 5],
            "all_ones": [1] * 20,
            "all_rm1": [C.r - 1] * 20,
            "single_nonzero": [0] * 15 + [rng.randrange(1, C.r)],
            "small_ascii": [ord(ch) for ch in "hello there my name is bob"],
        }
        openings = [(0, 1), (3, 1), (0, 4), (5, 7), (2, 9)]
        cases = []
        with tempfile.TemporaryDirectory(dir=os.path.join(ROOT, "oracle", "_ref")) as td:
            f = lambda n: os.path.join(td, n)
            for name, P in polys.items():
                open(f("c"), "wb").write(C.scalars_to_bytes(P))
                case = {"name": name, "coeffs": C.scalars_to_bytes(P).hex()}
                run(binary, "commit", srs, f("c"), f("o"))
                case["commit"] = open(f("o"), "rb").read().hex()
                run(binary, "polyser", f("c"), f("o"))
                case["polyser"] = open(f("o"), "rb").read().hex()
                run(binary, "g2msm", srs, f("c"), f("o"))
                case["g2msm"] = open(f("o"), "rb").read().hex()
                case["openings"] = []
                n = len(C.poly_trim(P))
                for off, k in openings + [(0, n), (1, max(n - 1, 1))]:
                    if k < 1:
                        continue
                    op = {"off": off, "k": k}
                    run(binary, "prove", srs, f("c"), off, k, f("o"))
                    op["proof"] = open(f("o"), "rb").read().hex()
                    run(binary, "quotient", f("c"), off, k, f("o"))
                    op["quotient"] = open(f("o"), "rb").read().hex()
                    run(binary, "eval", f("c"), off, k, f("y"))
                    op["ys"] = open(f("y"), "rb").read().hex()
                    run(binary, "interp", off, f("y"), f("o"))
                    op["interp"] = open(f("o"), "rb").read().hex()
                    run(binary, "vanishing", off, k, f("o"))
                    op["vanishing"] = open(f("o"), "rb").read().hex()
                    # reference verify_proof on its own commit/proof, and on tampered data
                    open(f("cm"), "wb").write(bytes.fromhex(case["commit"]))
                    open(f("pf"), "wb").write(bytes.fromhex(op["proof"]))
                    op["verify"] = int(run(binary, "verify", srs, f("cm"), f("pf"), off, f("y")).strip())
                    ys = bytearray(open(f("y"), "rb").read())
                    ys[0] ^= 1
                    open(f("y2"), "wb").write(bytes(ys))
                    op["verify_tampered"] = int(run(binary, "verify", srs, f("cm"), f("pf"), off, f("y2")).strip())
                    op["verify_wrong_offset"] = int(run(binary, "verify", srs, f("cm"), f("pf"), off + 1, f("y")).strip())
                    case["openings"].append(op)
                cases.append(case)
        json.dump({"curve": cname, "srs_n": SRS_N, "cases": cases},
                  open(os.path.join(outdir, "kat.json"), "w"), indent=0)
        print(cname, "cases:", len(cases), "size:", os.path.getsize(os.path.join(outdir, "kat.json")))


if __name__ == "__main__":
    main()
