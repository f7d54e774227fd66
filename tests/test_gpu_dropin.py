"""The C++ drop-in (src/trusted_setup_gpu.cpp in place of the reference's src/trusted_setup.cpp, everything
else of the reference unmodified; built by src/build_dropin.sh into oracle/_ref/dropin/) on a real GPU:

  * the reference's OWN test program (testing/testing.cpp, 36 property checks incl. the two EIP-4844 blobs)
    and its own benchmark (benchmark/benchmark.cpp) run unchanged on top of the CUDA path;
  * the byte-file harness (oracle/ref_harness.cpp) linked against the drop-in reproduces the golden vectors
    the unmodified reference produced (tests/golden), bit for bit, and its verify_proof (device MSMs + host
    pairings) accepts the golden proofs and rejects tampered data;
  * cross-verification: proofs made by the drop-in are accepted by the unmodified reference binary.
"""
import os
import re
import subprocess
import tempfile

import pytest

from tests.golden_util import CURVE_NAMES, ROOT, load

pytestmark = pytest.mark.gpu
DROPIN = os.path.join(ROOT, "oracle", "_ref", "dropin")
REF = os.path.join(ROOT, "oracle", "_ref")


def _need(path):
    if not os.path.exists(path):
        pytest.skip(f"{os.path.relpath(path, ROOT)} not built (src/build_dropin.sh needs /root/reference)")
    return path


def _run(binary, *args, cwd=None, timeout=600):
    return subprocess.run([binary, *map(str, args)], check=True, capture_output=True, text=True, cwd=cwd,
                          timeout=timeout).stdout


def test_reference_test_program_on_gpu():
    """testing/testing.cpp, unmodified, linked against the drop-in: every check must print PASSED."""
    out = _run(_need(os.path.join(DROPIN, "testing-bn254-gpu")), cwd=DROPIN)
    plain = re.sub(r"\x1b\[[0-9;]*m", "", out)
    assert "FAILED" not in plain, plain
    assert plain.count("PASSED") == 36, plain


def test_dropin_concurrent_callers_and_cache_eviction():
    """tests/cpp/dropin_threads.cpp: 8 threads making the first use of one file-loaded setup, 11 live setups (more than
    the resident-SRS cache holds) used from 4 threads, copies sharing one resident SRS."""
    with tempfile.TemporaryDirectory() as td:
        out = _run(_need(os.path.join(DROPIN, "dropin-threads-bn254")), os.path.join(td, "srs.bin"), cwd=DROPIN)
    assert out.strip().startswith("OK"), out


@pytest.mark.parametrize("cname", ["bn254", "bls12381"])
def test_reference_benchmark_on_gpu(cname):
    """benchmark/benchmark.cpp, unmodified: every row verifies (BASELINE config 1 / 5 shapes, on the GPU path)."""
    out = _run(_need(os.path.join(DROPIN, f"benchmark-{cname}-gpu")), cwd=DROPIN)
    assert "✗" not in out and "✓" in out, out
    outdir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(outdir):  # keep the table next to the other GPU-run artefacts
        with open(os.path.join(outdir, f"benchmark-{cname}-gpu.txt"), "w") as f:
            f.write(out)


@pytest.mark.parametrize("cname", CURVE_NAMES)
def test_dropin_reproduces_reference_golden_vectors(cname):
    drop = _need(os.path.join(DROPIN, f"kzgdrop-{cname}"))
    ref = os.path.join(REF, f"kzgref-{cname}")
    kat, srs_bytes = load(cname)
    want = {c["name"]: c for c in kat["cases"]}
    with tempfile.TemporaryDirectory() as td:
        f = lambda n: os.path.join(td, n)
        open(f("srs"), "wb").write(srs_bytes)
        for name in ("zero", "deg37", "deg46_max", "edge_scalars"):
            case = want[name]
            open(f("c"), "wb").write(bytes.fromhex(case["coeffs"]))
            _run(drop, "commit", f("srs"), f("c"), f("o"))
            assert open(f("o"), "rb").read().hex() == case["commit"], name
            _run(drop, "g2msm", f("srs"), f("c"), f("o"))
            assert open(f("o"), "rb").read().hex() == case["g2msm"], name
            for op in case["openings"][2:5:2]:
                off, k = op["off"], op["k"]
                _run(drop, "prove", f("srs"), f("c"), off, k, f("pf"))
                assert open(f("pf"), "rb").read().hex() == op["proof"], (name, off, k)
                open(f("cm"), "wb").write(bytes.fromhex(case["commit"]))
                open(f("y"), "wb").write(bytes.fromhex(op["ys"]))
                # the drop-in's verify_proof: device MSMs over Z and I, host pairings
                assert int(_run(drop, "verify", f("srs"), f("cm"), f("pf"), off, f("y")).strip()) == op["verify"]
                bad = bytearray(bytes.fromhex(op["ys"]))
                bad[0] ^= 1
                open(f("y2"), "wb").write(bytes(bad))
                assert int(_run(drop, "verify", f("srs"), f("cm"), f("pf"), off, f("y2")).strip()) == op["verify_tampered"]
                assert int(_run(drop, "verify", f("srs"), f("cm"), f("pf"), off + 1, f("y")).strip()) == op["verify_wrong_offset"]
                if os.path.exists(ref):  # the unmodified reference accepts the drop-in's proof
                    assert int(_run(ref, "verify", f("srs"), f("cm"), f("pf"), off, f("y")).strip()) == op["verify"]


def test_dropin_trusted_setup_is_loadable_by_reference():
    """trusted_setup(int) on the device -> export_setup -> the UNMODIFIED reference loads the file, commits to
    a polynomial with it and gets the same bytes as the drop-in."""
    cname = "bls12381"
    drop = _need(os.path.join(DROPIN, f"kzgdrop-{cname}"))
    ref = _need(os.path.join(REF, f"kzgref-{cname}"))
    with tempfile.TemporaryDirectory() as td:
        f = lambda n: os.path.join(td, n)
        _run(drop, "setup", 40, f("srs"))
        coeffs = bytes(range(1, 33)) * 30
        coeffs = b"".join(coeffs[i:i + 31] + b"\x00" for i in range(0, 31 * 30, 31))
        open(f("c"), "wb").write(coeffs)
        _run(drop, "commit", f("srs"), f("c"), f("o1"))
        _run(ref, "commit", f("srs"), f("c"), f("o2"))
        assert open(f("o1"), "rb").read() == open(f("o2"), "rb").read()
        _run(drop, "prove", f("srs"), f("c"), 2, 5, f("p1"))
        _run(ref, "prove", f("srs"), f("c"), 2, 5, f("p2"))
        assert open(f("p1"), "rb").read() == open(f("p2"), "rb").read()


def test_demo_cli_flow_cross_process():
    """the reference's demo CLI (demo/shared/kzg-cli.cpp, unmodified) linked against the drop-in,
    driven the way demo/run-demo drives it — setup 65536, commit a 142 KB document (4583 chunks of 31 bytes),
    random 4-chunk subsection proofs, verify — with the UNMODIFIED reference CLI as the other party: it loads
    the GPU-made kzg_public file, must produce the same commitment and proof hex, and must accept / reject
    the same transcripts."""
    import random
    gpu = _need(os.path.join(DROPIN, "kzg-cli-bn254-gpu"))
    ref = _need(os.path.join(REF, "kzg-cli-bn254"))
    rng = random.Random(142054)
    words = ["kzg", "commit", "proof", "ledger", "peer", "host", "block", "chunk", "pairing", "setup", "\n", "blob"]
    doc = ""
    while len(doc) < 142054:
        doc += rng.choice(words) + " "
    doc = doc[:142054]
    with tempfile.TemporaryDirectory() as td:
        shared, peer = os.path.join(td, "shared"), os.path.join(td, "peer-a")
        os.mkdir(shared)
        os.mkdir(peer)
        with open(os.path.join(peer, "document.txt"), "w") as f:
            f.write(doc)
        out = _run(gpu, "setup", 65536, cwd=shared)
        assert "num_coeff=65536" in out and os.path.getsize(os.path.join(shared, "kzg_public")) > 65536 * 3 * 32
        commit = _run(gpu, "commit", "document.txt", cwd=peer).strip()
        assert re.fullmatch(r"[0-9a-f]+", commit)
        assert _run(ref, "commit", "document.txt", cwd=peer, timeout=1200).strip() == commit
        # the same document through the C ABI's fused device chain (kzg_gpu_commit_bytes), on the CLI-made setup file
        import sys
        sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
        import kzg_b200
        with open(os.path.join(shared, "kzg_public"), "rb") as f:
            ts = kzg_b200.TrustedSetup.from_file_bytes("bn254", f.read())
        raw = doc.encode("latin-1")
        raw += bytes(31 - len(raw) % 31)  # kzg-cli.cpp:50-54
        fused, _ = ts.commit_bytes(raw, 31)
        assert fused.serialize().hex() == commit
        ts.ctx.close()
        for seed in (7, 123456789):
            proof, offset, data = _run(gpu, "prove", "document.txt", seed, cwd=peer).split()
            if seed == 7:  # the reference prover takes seconds per proof: compare one
                assert _run(ref, "prove", "document.txt", seed, cwd=peer, timeout=1200).split() == [proof, offset, data]
            for cli in (gpu, ref):
                ok = subprocess.run([cli, "verify", commit, proof, offset, data], cwd=peer, capture_output=True, timeout=1200)
                assert ok.returncode == 0, (cli, ok.stderr)
                bad = data[:10] + ("0" if data[10] != "0" else "1") + data[11:]
                no = subprocess.run([cli, "verify", commit, proof, offset, bad], cwd=peer, capture_output=True, timeout=1200)
                assert no.returncode == 1, (cli, no.stderr)
