#!/usr/bin/env python3
"""Known answers AT BASELINE SIZES from the UNMODIFIED reference (oracle/_ref/kzgref-<curve>), SURVEY.md 8(c)
last paragraph: degree {149, 4096} polynomials, k in {1, 4, 57, 128, 4096}, the reference's own fixtures
testing/blob1.txt / blob2.txt chunked as eth_blob_test does (testing/testing.cpp:53-102), the strings of
high_poly_degree_test (testing/testing.cpp:222-252), and config 5 of BASELINE.json as SURVEY 8(d) states it:
BN254, a degree-8191 polynomial opened at k = 4096 points so that the quotient is not zero.

The SRS must be the same on both sides and the reference draws its secret at random, so the SRS of this
fixture set is DERIVED: [s^i]G1, [s^i]G2 for a secret s written in the fixture, computed here by the CPU oracle
(oracle/kzg_oracle.py setup(), a few minutes of big-int arithmetic spread over the host cores), written in the
reference's file format and LOADED BY THE REFERENCE (its loader checks every point is on the curve).  Only the
sha256 of that file is committed: the GPU tests regenerate the SRS on the device from s (kzg_gpu_setup_powers),
require the same sha256, and must then reproduce the reference's commit.serialize() / proof.serialize() /
polyeval_G2 bytes.  Polynomials are committed as recipes (seeded sha256 counter stream, or the fixture file they
come from) plus the sha256 of the coefficient bytes the REFERENCE's own polyfit produced.

Run in the container that has /root/reference:  python tools/gen_golden_large.py      (about 10 minutes)
Writes tests/golden/<curve>/kat_large.json.
"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile
from multiprocessing import Pool

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.kzg_oracle import get_curve  # noqa: E402
from tests.golden_util import chunk_values, ref_from_hex, stream_scalars  # noqa: E402

REF_TESTING = "/root/reference/testing"
STR_149 = ("wrgJKdE3t5bECALy3eKIwYxEF3V7Z8KTx0nFe1IX5tjH22F5gXOa5LnIMIQuOiNJj8YL8rqDiZSkZfoEDAmGTXXqqvkCd5WKE2fMtVXa2zKae6opGY4i6bY"
           "uUG67LaSXd5tUbO4bNPB0TxnkWrSaQ")  # testing/testing.cpp:233 (150 characters -> degree 149)


def _setup_chunk(args):
    cname, s, lo, hi = args
    C = get_curve(cname)
    g1, g2 = [], []
    si = pow(s, lo, C.r)
    for _ in range(lo, hi):
        g1.append(C.mul(si, C.g1))
        g2.append(C.mul(si, C.g2, g2=True))
        si = si * s % C.r
    return lo, g1, g2


def make_srs(cname, s, n, path):
    C = get_curve(cname)
    step = 64
    jobs = [(cname, s, lo, min(n, lo + step)) for lo in range(0, n, step)]
    with Pool(os.cpu_count() or 4) as pool:
        parts = sorted(pool.map(_setup_chunk, jobs))
    g1 = [p for _, a, _ in parts for p in a]
    g2 = [p for _, _, b in parts for p in b]
    raw = C.srs_to_file_bytes(g1, g2)
    open(path, "wb").write(raw)
    return hashlib.sha256(raw).hexdigest()


def run(binary, *args):
    return subprocess.run([binary, *map(str, args)], check=True, capture_output=True, text=True).stdout


def main():
    plan = {
        "bls12381": {"n": 5000, "secret": 0x4B5A47_5EC2E7_0F_B15_12381_0123456789ABCDEF_0FEDCBA987654321 % get_curve("bls12381").r,
                     "big": None},
        "bn254": {"n": 8200, "secret": 0x4B5A47_5EC2E7_0F_B2_254_0123456789ABCDEF_0FEDCBA987654321 % get_curve("bn254").r,
                  "big": 8192},
    }
    only = [a for a in sys.argv[1:] if not a.startswith("--")] or list(plan)
    for cname in only:
        cfg = plan[cname]
        C = get_curve(cname)
        binary = os.path.join(ROOT, "oracle", "_ref", f"kzgref-{cname}")
        outdir = os.path.join(ROOT, "tests", "golden", cname)
        work = os.path.join(ROOT, "oracle", "_ref", "large")
        os.makedirs(work, exist_ok=True)
        srs = os.path.join(work, f"srs-{cname}.bin")
        if os.path.exists(srs) and "--fresh" not in sys.argv:
            sha = hashlib.sha256(open(srs, "rb").read()).hexdigest()
        else:
            sha = make_srs(cname, cfg["secret"], cfg["n"], srs)
        print(cname, "srs", cfg["n"], sha, flush=True)
        chunk = C.SB - 1
        polys = []  # (name, recipe, coefficient source)
        polys.append(("deg149_string", {"kind": "string", "text": STR_149, "offset": 0}, ("ys", [ord(c) for c in STR_149])))
        polys.append(("deg4096_uniform", {"kind": "stream", "seed": "deg4096-" + cname, "n": 4097},
                      ("coeffs", stream_scalars(C.r, "deg4096-" + cname, 4097))))
        for bname in ("blob1", "blob2"):
            data = ref_from_hex(open(os.path.join(REF_TESTING, bname + ".txt")).read())
            ys = chunk_values(data, chunk)
            if len(ys) + 1 < cfg["n"]:
                polys.append((bname, {"kind": "blob_file", "file": bname + ".txt", "chunk": chunk, "sha256_bytes":
                                      hashlib.sha256(data).hexdigest(), "chunks": len(ys)}, ("ys", ys)))
        if cfg["big"]:
            nb = cfg["big"]
            polys.append(("deg%d_uniform" % (nb - 1), {"kind": "stream", "seed": "big-" + cname, "n": nb},
                          ("coeffs", stream_scalars(C.r, "big-" + cname, nb))))
        cases = []
        with tempfile.TemporaryDirectory(dir=work) as td:
            f = lambda nme: os.path.join(td, nme)
            for name, recipe, (kind, vals) in polys:
                if kind == "ys":   # the reference interpolates (polyfit through consecutive nodes 0..k-1)
                    open(f("y"), "wb").write(C.scalars_to_bytes(vals))
                    run(binary, "interp", 0, f("y"), f("c"))
                    coeffs_raw = open(f("c"), "rb").read()
                else:
                    coeffs_raw = C.scalars_to_bytes(vals)
                    open(f("c"), "wb").write(coeffs_raw)
                n = len(C.poly_trim(C.scalars_from_bytes(coeffs_raw)))
                case = {"name": name, "recipe": recipe, "n_coeffs": n, "coeffs_sha256": hashlib.sha256(coeffs_raw[:n * C.SB]).hexdigest()}
                run(binary, "commit", srs, f("c"), f("o"))
                case["commit"] = open(f("o"), "rb").read().hex()
                ks = [1, 4, 57, 128, 4096]
                offs = {1: 17, 4: 1000 if n > 1100 else 3, 57: 49, 128: 0, 4096: 0}
                case["openings"] = []
                for k in ks:
                    if k >= n + 64:
                        continue
                    off = offs[k]
                    op = {"off": off, "k": k}
                    run(binary, "prove", srs, f("c"), off, k, f("o"))
                    op["proof"] = open(f("o"), "rb").read().hex()
                    run(binary, "eval", f("c"), off, k, f("yy"))
                    ysb = open(f("yy"), "rb").read()
                    op["ys_sha256"] = hashlib.sha256(ysb).hexdigest()
                    open(f("cm"), "wb").write(bytes.fromhex(case["commit"]))
                    open(f("pf"), "wb").write(bytes.fromhex(op["proof"]))
                    if k < cfg["n"]:
                        op["verify"] = int(run(binary, "verify", srs, f("cm"), f("pf"), off, f("yy")).strip())
                        t = bytearray(ysb); t[len(t) // 2] ^= 0x10
                        open(f("y2"), "wb").write(bytes(t))
                        op["verify_tampered"] = int(run(binary, "verify", srs, f("cm"), f("pf"), off, f("y2")).strip())
                        op["verify_wrong_offset"] = int(run(binary, "verify", srs, f("cm"), f("pf"), off + 1, f("yy")).strip())
                        if k > 1:
                            open(f("y3"), "wb").write(ysb[:-C.SB])
                            op["verify_short"] = int(run(binary, "verify", srs, f("cm"), f("pf"), off, f("y3")).strip())
                    if k >= 57:   # verify-side MSMs on the reference's own Z (polyeval_G2) — config 5's G2 MSM
                        run(binary, "vanishing", off, k, f("z"))
                        op["vanishing_sha256"] = hashlib.sha256(open(f("z"), "rb").read()).hexdigest()
                        run(binary, "g2msm", srs, f("z"), f("o"))
                        op["g2msm_vanishing"] = open(f("o"), "rb").read().hex()
                    case["openings"].append(op)
                    print(cname, name, "k", k, {a: b for a, b in op.items() if a.startswith("verify")}, flush=True)
                cases.append(case)
        json.dump({"curve": cname, "srs_n": cfg["n"], "secret": hex(cfg["secret"]), "srs_sha256": sha,
                   "generator": "tools/gen_golden_large.py", "cases": cases},
                  open(os.path.join(outdir, "kat_large.json"), "w"), indent=1)
        print(cname, "done", flush=True)


if __name__ == "__main__":
    main()
