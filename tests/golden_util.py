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


# ---- recipes shared with tools/gen_golden_large.py (the large fixtures store recipes, not coefficient bytes) ----
import hashlib  # noqa: E402
import re  # noqa: E402


def load_large(curve):
    return json.load(open(os.path.join(ROOT, "tests", "golden", curve, "kat_large.json")))


def stream_scalars(r, seed, n):
    """c_i = sha256(seed | i) (taken twice, 64 bytes) mod r."""
    return [int.from_bytes(hashlib.sha256(b"%s|%d" % (seed.encode(), i)).digest() * 2, "big") % r for i in range(n)]


def ref_from_hex(text):
    """testing/testing.cpp:406-413 from_hex, faithfully: two characters at a time through strtol(.., 16), so a
    trailing newline contributes a zero byte (blob1.txt) and a dangling single digit is a byte of its own (blob2.txt)."""
    out = bytearray()
    for i in range(0, len(text), 2):
        m = re.match(r"\s*([0-9a-fA-F]*)", text[i:i + 2])
        out.append(int(m.group(1), 16) & 0xFF if m.group(1) else 0)
    return bytes(out)


def chunk_values(data, chunk):
    """eth_blob_test's zero padding (a FULL chunk when already aligned, testing.cpp:63-65) + blob::from_bytes chunks."""
    data = data + bytes(chunk - (len(data) % chunk))
    return [int.from_bytes(data[i:i + chunk], "little") for i in range(0, len(data), chunk)]
