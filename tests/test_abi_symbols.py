"""CPU-side checks of the C-ABI boundary: libkzg_gpu.so loads and exports every function include/kzg_gpu.h declares,
the Python mirror binds every one of them, and (no GPU here) creating a context fails loudly instead of falling back
to a CPU path.  No compute call is made."""
import ctypes
import os
import re
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import kzg_b200  # noqa: E402


def declared_functions():
    text = open(os.path.join(ROOT, "include", "kzg_gpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(kzg_gpu_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = declared_functions()
    assert len(names) >= 40
    lib = ctypes.CDLL(kzg_b200.LIB_PATH)
    for name in names:
        assert getattr(lib, name) is not None, name


def test_python_mirror_binds_every_declared_symbol():
    assert sorted(kzg_b200.EXPORTED_SYMBOLS) == declared_functions()
    lib = kzg_b200.load_library()
    for name in kzg_b200.EXPORTED_SYMBOLS:
        assert getattr(lib, name).argtypes is not None, name


def test_no_cpu_fallback_without_a_device():
    lib = kzg_b200.load_library()
    if lib.kzg_gpu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(RuntimeError):
        kzg_b200.Context("bls12381", 0)
    with pytest.raises(RuntimeError):
        kzg_b200.Group("bls12381")


def test_bench_host_side_checkers_agree_with_the_oracle():
    """bench.py's post-run parity check evaluates sum_i c_i s^(first + i) and per-coefficient sums over blobs on the host."""
    import numpy as np
    sys.path.insert(0, ROOT)
    import bench
    from oracle.kzg_oracle import get_curve
    C = get_curve("bls12381")
    arr = bench.random_scalars(24, 3)
    ints = bench.scalars_to_ints(arr)
    assert all(0 <= v < C.r for v in ints)
    assert ints == C.scalars_from_bytes(arr.view(np.uint8).tobytes())
    s = 0x1234567
    assert bench.fr_eval(ints, s, C.r) == C.poly_eval(ints, s)
    assert bench.fr_eval(ints, s, C.r, first=5) == C.poly_eval([0] * 5 + ints, s)
    cols = bench.column_sums(arr, 4, 6)
    assert cols == [sum(ints[b * 6 + i] for b in range(4)) for i in range(6)]
