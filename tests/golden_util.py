"""Loader for tests/golden/<curve>/ (written by tools/gen_golden.py from the real reference)."""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CURVE_NAMES = ("bls12381", "bn254", "bn158")


def load(curve):
    d = os.path.join(ROOT, "tests", "golden", curve)
    kat = json.load(open(os.path.join(d, "kat.json")))
    srs = open(os.path.join(d, "srs.bin"), "rb").read()
    return kat, srs
