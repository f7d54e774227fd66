"""Host-side mirror of the reference's kzg.h API (src/kzg.h:27-292) over the CUDA C ABI (include/kzg_gpu.h).

The reference's host language is C++ (the C++ drop-in for src/trusted_setup.cpp etc. lives in src/ of this
repository and needs MIRACL + NTL to build); this module is the same boundary for Python callers, tests and
bench.py.  Names and argument meaning follow kzg.h:

    kzg::trusted_setup(int)            -> TrustedSetup.generate(curve, n, secret)      src/trusted_setup.cpp:21-74
    kzg::trusted_setup(const string&)  -> TrustedSetup.from_file_bytes(curve, raw)     src/trusted_setup.cpp:76-121
    export_setup                       -> TrustedSetup.export_setup()                  src/trusted_setup.cpp:256-287
    create_commit / verify_commit      -> same names                                   src/trusted_setup.cpp:137-147
    create_proof(poly, off, len)       -> create_proof                                 src/trusted_setup.cpp:214-228
    verify_proof: its two MSMs         -> verify_proof_msms (polyeval_G2(Z), polyeval_G1(I))   :239-248
    kzg::poly / commit / proof         -> Poly / Commit / Proof with serialize()/deserialize() (src/util.cpp:78-170)

Error behaviour matches the reference: the argument checks raise ValueError where the reference throws
std::invalid_argument, with the same messages; a failing CUDA call raises RuntimeError.  There is NO CPU
fallback: importing works everywhere, but creating a Context without the CUDA library or a GPU raises.
The final pairing check of verify_proof stays in the reference's host code (MIRACL PAIR_ate/PAIR_fexp,
src/trusted_setup.cpp:242-253) and is therefore not part of this module.
"""
import ctypes
import os
import struct

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkzg_gpu.so")

CURVE_IDS = {"bn158": 0, "bn254": 1, "bls12381": 2}
# group order r per curve (CURVE_Order of the MIRACL ROM files; see csrc/curve_constants.cuh)
CURVE_ORDER = {
    0: 0x24012027628c9953cf6f73fa14d59f209f04200d,
    1: 0x2523648240000001ba344d8000000007ff9f800000000010a10000000000000d,
    2: 0x73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001,
}

_lib = None


def load_library():
    """dlopen libkzg_gpu.so (built in-tree by `make -C kzg-commitments_b200`).  Fails loudly when absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `make -C kzg-commitments_b200 -j4` "
                           "(there is no CPU fallback for the KZG hot path)")
    lib = ctypes.CDLL(LIB_PATH)
    u8p, u64, i64, ip, vp = ctypes.c_char_p, ctypes.c_uint64, ctypes.c_int64, ctypes.POINTER(ctypes.c_int), ctypes.c_void_p
    sigs = {
        "kzg_gpu_strerror": (ctypes.c_char_p, [ctypes.c_int]),
        "kzg_gpu_last_error": (ctypes.c_char_p, [vp]),
        "kzg_gpu_modbytes": (ctypes.c_int, [ctypes.c_int]),
        "kzg_gpu_scalar_bytes": (ctypes.c_int, [ctypes.c_int]),
        "kzg_gpu_ctx_create": (ctypes.c_int, [ctypes.POINTER(vp), ctypes.c_int, ctypes.c_int]),
        "kzg_gpu_ctx_destroy": (None, [vp]),
        "kzg_gpu_set_option": (ctypes.c_int, [vp, ctypes.c_char_p, ctypes.c_longlong]),
        "kzg_gpu_srs_load": (ctypes.c_int, [vp, u64, u8p, u8p]),
        "kzg_gpu_setup_powers": (ctypes.c_int, [vp, u8p, u64, u8p, u8p]),
        "kzg_gpu_setup_powers_range": (ctypes.c_int, [vp, u8p, u64, u64, u8p, u8p]),
        "kzg_gpu_msm_g1": (ctypes.c_int, [vp, vp, u64, u64, u8p, ip]),
        "kzg_gpu_msm_g2": (ctypes.c_int, [vp, vp, u64, u64, u8p, ip]),
        "kzg_gpu_msm_g1_batch": (ctypes.c_int, [vp, vp, u64, u64, u8p, ip]),
        "kzg_gpu_msm_g1_launch": (ctypes.c_int, [vp, vp, u64, u64, u64]),
        "kzg_gpu_msm_fetch": (ctypes.c_int, [vp, u8p, ip]),
        "kzg_gpu_msm_fetch_partial": (ctypes.c_int, [vp, ctypes.POINTER(ctypes.c_uint32), ctypes.POINTER(ctypes.c_size_t)]),
        "kzg_gpu_g1_sum_partials": (ctypes.c_int, [vp, ctypes.POINTER(ctypes.c_uint32), u64, u8p, ip]),
        "kzg_gpu_msm_partial_device": (ctypes.c_int, [vp, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]),
        "kzg_gpu_g1_sum_partials_device": (ctypes.c_int, [vp, vp, u64, u8p, ip]),
        "kzg_gpu_poly_quotient": (ctypes.c_int, [vp, u8p, u64, i64, u64, u8p]),
        "kzg_gpu_poly_eval_range": (ctypes.c_int, [vp, u8p, u64, i64, u64, u8p]),
        "kzg_gpu_poly_eval_range_noalloc": (ctypes.c_int, [vp, vp, u64, i64, u64, vp]),
        "kzg_gpu_prove": (ctypes.c_int, [vp, u8p, u64, i64, u64, u8p, ip]),
        "kzg_gpu_interpolate_range": (ctypes.c_int, [vp, u8p, i64, u64, u8p, u8p]),
        "kzg_gpu_commit_bytes": (ctypes.c_int, [vp, vp, u64, ctypes.c_int, i64, u8p, u8p, ctypes.POINTER(ctypes.c_int)]),
        "kzg_gpu_prove_batch": (ctypes.c_int, [vp, vp, u64, ctypes.POINTER(ctypes.c_int64), u64, u64, u8p, ip]),
        "kzg_gpu_device_count": (ctypes.c_int, []),
        "kzg_gpu_group_create": (ctypes.c_int, [ctypes.POINTER(vp), ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.c_int]),
        "kzg_gpu_group_destroy": (None, [vp]),
        "kzg_gpu_group_size": (ctypes.c_int, [vp]),
        "kzg_gpu_group_ctx": (vp, [vp, ctypes.c_int]),
        "kzg_gpu_group_last_error": (ctypes.c_char_p, [vp]),
        "kzg_gpu_group_set_option": (ctypes.c_int, [vp, ctypes.c_char_p, ctypes.c_longlong]),
        "kzg_gpu_group_srs_load": (ctypes.c_int, [vp, u64, u8p, u8p]),
        "kzg_gpu_group_setup_powers": (ctypes.c_int, [vp, u8p, u64, u8p, u8p]),
        "kzg_gpu_group_msm_g1": (ctypes.c_int, [vp, vp, u64, u64, u8p, ip]),
        "kzg_gpu_group_msm_g1_batch": (ctypes.c_int, [vp, vp, u64, u64, u8p, ip]),
        "kzg_gpu_group_prove_batch": (ctypes.c_int, [vp, vp, u64, ctypes.POINTER(ctypes.c_int64), u64, u64, u8p, ip]),
        "kzg_gpu_ctx_stream": (vp, [vp]),
        "kzg_gpu_launch_count": (u64, [vp]),
        "kzg_gpu_microbench": (ctypes.c_int, [vp, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]),
        "kzg_gpu_selftest": (ctypes.c_int, [vp, ip]),
        "kzg_gpu_phase_times": (ctypes.c_int, [vp, ctypes.POINTER(ctypes.c_float), ctypes.c_int]),
        "kzg_gpu_msm_stats": (ctypes.c_int, [vp, ctypes.POINTER(ctypes.c_uint64), ctypes.c_int]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export the declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


EXPORTED_SYMBOLS = [
    "kzg_gpu_strerror", "kzg_gpu_last_error", "kzg_gpu_modbytes", "kzg_gpu_scalar_bytes", "kzg_gpu_ctx_create",
    "kzg_gpu_ctx_destroy", "kzg_gpu_set_option", "kzg_gpu_srs_load", "kzg_gpu_setup_powers", "kzg_gpu_msm_g1",
    "kzg_gpu_msm_g2", "kzg_gpu_msm_g1_batch", "kzg_gpu_msm_g1_launch", "kzg_gpu_msm_fetch",
    "kzg_gpu_msm_fetch_partial", "kzg_gpu_g1_sum_partials", "kzg_gpu_poly_quotient", "kzg_gpu_poly_eval_range",
    "kzg_gpu_prove", "kzg_gpu_ctx_stream", "kzg_gpu_launch_count", "kzg_gpu_microbench", "kzg_gpu_selftest",
    "kzg_gpu_phase_times", "kzg_gpu_msm_stats", "kzg_gpu_msm_partial_device", "kzg_gpu_g1_sum_partials_device",
    "kzg_gpu_interpolate_range", "kzg_gpu_prove_batch", "kzg_gpu_commit_bytes", "kzg_gpu_setup_powers_range",
    "kzg_gpu_poly_eval_range_noalloc",
    "kzg_gpu_device_count", "kzg_gpu_group_create", "kzg_gpu_group_destroy", "kzg_gpu_group_size", "kzg_gpu_group_ctx",
    "kzg_gpu_group_last_error", "kzg_gpu_group_set_option", "kzg_gpu_group_srs_load", "kzg_gpu_group_setup_powers",
    "kzg_gpu_group_msm_g1", "kzg_gpu_group_msm_g1_batch", "kzg_gpu_group_prove_batch",
]


def _ptr(buf):
    """host bytes / bytearray / ctypes buffer / integer address -> c_void_p"""
    if isinstance(buf, int):
        return ctypes.c_void_p(buf)
    if isinstance(buf, bytes):
        return ctypes.cast(ctypes.c_char_p(buf), ctypes.c_void_p)
    if isinstance(buf, bytearray):
        return ctypes.cast((ctypes.c_char * len(buf)).from_buffer(buf), ctypes.c_void_p)
    if hasattr(buf, "data_ptr"):      # torch tensor (pinned host or device memory)
        return ctypes.c_void_p(buf.data_ptr())
    if hasattr(buf, "ctypes"):        # numpy array
        return ctypes.c_void_p(buf.ctypes.data)
    return ctypes.cast(buf, ctypes.c_void_p)


class Context:
    """One kzg_gpu_ctx: a curve parameter set on one CUDA device, with a resident SRS."""

    def __init__(self, curve, device=0):
        self.lib = load_library()
        self.curve = CURVE_IDS[curve.lower()] if isinstance(curve, str) else int(curve)
        self.M = self.lib.kzg_gpu_modbytes(self.curve)
        self.SB = self.lib.kzg_gpu_scalar_bytes(self.curve)
        if self.M < 0:
            raise ValueError("unknown curve")
        self.r = CURVE_ORDER[self.curve]
        h = ctypes.c_void_p()
        rc = self.lib.kzg_gpu_ctx_create(ctypes.byref(h), self.curve, device)
        if rc != 0:
            raise RuntimeError("kzg_gpu_ctx_create failed: %s (a CUDA device is required; no CPU fallback)"
                               % self.lib.kzg_gpu_strerror(rc).decode())
        self.h = h
        self.srs_n = 0

    def close(self):
        if getattr(self, "h", None):
            self.lib.kzg_gpu_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError("%s failed: %s [%s]" % (what, self.lib.kzg_gpu_strerror(rc).decode(),
                                                       self.lib.kzg_gpu_last_error(self.h).decode()))

    def set_option(self, name, value):
        self._check(self.lib.kzg_gpu_set_option(self.h, name.encode(), int(value)), "set_option " + name)

    def srs_load(self, g1, g2=None):
        n = len(g1) // (2 * self.M)
        if len(g1) != n * 2 * self.M or (g2 is not None and len(g2) != n * 4 * self.M):
            raise ValueError("SRS byte length does not match the curve's point size")
        self._check(self.lib.kzg_gpu_srs_load(self.h, n, g1, g2), "srs_load")
        self.srs_n = n

    def setup_powers(self, s, n, want_bytes=True, first_index=0):
        """[s^(first_index + i)]G1 / G2, i < n, left resident (first_index > 0: one device's slice of a shared SRS)."""
        s_le = (int(s) % self.r).to_bytes(self.SB, "little")
        g1 = ctypes.create_string_buffer(n * 2 * self.M) if want_bytes else None
        g2 = ctypes.create_string_buffer(n * 4 * self.M) if want_bytes else None
        self._check(self.lib.kzg_gpu_setup_powers_range(self.h, s_le, first_index, n, g1, g2), "setup_powers")
        self.srs_n = n
        return (g1.raw, g2.raw) if want_bytes else (None, None)

    def _msm(self, fn, scalars, count, base_offset, ncoord):
        out = ctypes.create_string_buffer(ncoord * self.M)
        inf = ctypes.c_int(0)
        self._check(fn(self.h, _ptr(scalars), count, base_offset, out, ctypes.byref(inf)), "msm")
        return out.raw, bool(inf.value)

    def msm_g1(self, scalars, count=None, base_offset=0):
        count = len(scalars) // self.SB if count is None else count
        return self._msm(self.lib.kzg_gpu_msm_g1, scalars, count, base_offset, 2)

    def msm_g2(self, scalars, count=None, base_offset=0):
        count = len(scalars) // self.SB if count is None else count
        return self._msm(self.lib.kzg_gpu_msm_g2, scalars, count, base_offset, 4)

    def msm_g1_batch(self, scalars, count, batch):
        out = ctypes.create_string_buffer(batch * 2 * self.M)
        inf = (ctypes.c_int * batch)()
        self._check(self.lib.kzg_gpu_msm_g1_batch(self.h, _ptr(scalars), count, batch, out, inf), "msm_g1_batch")
        return out.raw, [bool(v) for v in inf]

    def msm_g1_launch(self, d_scalars, count, batch=1, base_offset=0):
        self._check(self.lib.kzg_gpu_msm_g1_launch(self.h, _ptr(d_scalars), count, batch, base_offset), "msm_g1_launch")

    def msm_fetch(self, batch=1, ncoord=2):
        out = ctypes.create_string_buffer(batch * ncoord * self.M)
        inf = (ctypes.c_int * batch)()
        self._check(self.lib.kzg_gpu_msm_fetch(self.h, out, inf), "msm_fetch")
        return out.raw, [bool(v) for v in inf]

    def msm_fetch_partial(self):
        nw = ctypes.c_size_t(0)
        self._check(self.lib.kzg_gpu_msm_fetch_partial(self.h, None, ctypes.byref(nw)), "fetch_partial")
        words = (ctypes.c_uint32 * nw.value)()
        self._check(self.lib.kzg_gpu_msm_fetch_partial(self.h, words, ctypes.byref(nw)), "fetch_partial")
        return bytes(words)

    def msm_partial_device(self):
        """(device address, bytes) of the projective result of the most recent single G1 launch."""
        p, nb = ctypes.c_void_p(), ctypes.c_size_t(0)
        self._check(self.lib.kzg_gpu_msm_partial_device(self.h, ctypes.byref(p), ctypes.byref(nb)), "msm_partial_device")
        return p.value, nb.value

    def g1_sum_partials_device(self, d_ptr, n):
        out = ctypes.create_string_buffer(2 * self.M)
        inf = ctypes.c_int(0)
        self._check(self.lib.kzg_gpu_g1_sum_partials_device(self.h, _ptr(d_ptr), n, out, ctypes.byref(inf)), "g1_sum_partials_device")
        return out.raw, bool(inf.value)

    def g1_sum_partials(self, partials):
        raw = b"".join(partials)
        n = len(partials)
        words = (ctypes.c_uint32 * (len(raw) // 4)).from_buffer_copy(raw)
        out = ctypes.create_string_buffer(2 * self.M)
        inf = ctypes.c_int(0)
        self._check(self.lib.kzg_gpu_g1_sum_partials(self.h, words, n, out, ctypes.byref(inf)), "g1_sum_partials")
        return out.raw, bool(inf.value)

    def poly_quotient(self, coeffs, first_root, k):
        n = len(coeffs) // self.SB
        out = ctypes.create_string_buffer(max(n - k, 0) * self.SB)
        self._check(self.lib.kzg_gpu_poly_quotient(self.h, coeffs, n, first_root, k, out), "poly_quotient")
        return out.raw

    def poly_eval_range(self, coeffs, first_root, k):
        n = len(coeffs) // self.SB
        out = ctypes.create_string_buffer(k * self.SB)
        self._check(self.lib.kzg_gpu_poly_eval_range(self.h, coeffs, n, first_root, k, out), "poly_eval_range")
        return out.raw

    def poly_eval_range_noalloc(self, d_coeffs, n, first_root, k, d_y_out):
        """kzg_gpu_poly_eval_range_noalloc: DEVICE buffers in and out (tensors or addresses), nothing allocated, enqueued
        on the context's stream without waiting."""
        self._check(self.lib.kzg_gpu_poly_eval_range_noalloc(self.h, _ptr(d_coeffs), n, first_root, k, _ptr(d_y_out)),
                    "poly_eval_range_noalloc")

    def interpolate_range(self, ys, first_x, want_vanishing=True):
        """(I, Z) through the nodes first_x + i: I as k coefficient records, Z as k + 1 (or None)."""
        k = len(ys) // self.SB
        out = ctypes.create_string_buffer(k * self.SB)
        van = ctypes.create_string_buffer((k + 1) * self.SB) if want_vanishing else None
        self._check(self.lib.kzg_gpu_interpolate_range(self.h, ys, first_x, k, out, van), "interpolate_range")
        return out.raw, (van.raw if want_vanishing else None)

    def commit_bytes(self, data, chunk_size, chunk_offset=0, want_coeffs=True):
        """blob::from_bytes + poly::from_blob + create_commit on the device: (xy, is_inf, coefficient bytes or None)."""
        k = len(data) // chunk_size if chunk_size > 0 else 0
        out = ctypes.create_string_buffer(2 * self.M)
        co = ctypes.create_string_buffer(k * self.SB) if want_coeffs and k else None
        inf = ctypes.c_int(0)
        self._check(self.lib.kzg_gpu_commit_bytes(self.h, _ptr(data), len(data), chunk_size, chunk_offset, co, out,
                                                  ctypes.byref(inf)), "commit_bytes")
        return out.raw, bool(inf.value), (co.raw if co is not None else None)

    def prove(self, coeffs, first_root, k):
        n = len(coeffs) // self.SB
        out = ctypes.create_string_buffer(2 * self.M)
        inf = ctypes.c_int(0)
        self._check(self.lib.kzg_gpu_prove(self.h, coeffs, n, first_root, k, out, ctypes.byref(inf)), "prove")
        return out.raw, bool(inf.value)

    def prove_batch(self, coeffs, n, first_roots, k):
        batch = len(first_roots)
        out = ctypes.create_string_buffer(batch * 2 * self.M)
        inf = (ctypes.c_int * batch)()
        roots = (ctypes.c_int64 * batch)(*first_roots)
        self._check(self.lib.kzg_gpu_prove_batch(self.h, _ptr(coeffs), n, roots, k, batch, out, inf), "prove_batch")
        return out.raw, [bool(v) for v in inf]

    @property
    def stream(self):
        return self.lib.kzg_gpu_ctx_stream(self.h)

    @property
    def launch_count(self):
        return int(self.lib.kzg_gpu_launch_count(self.h))

    def microbench(self, kind):
        v = ctypes.c_double(0)
        self._check(self.lib.kzg_gpu_microbench(self.h, kind, ctypes.byref(v)), "microbench")
        return v.value

    def phase_times(self):
        """ms of the last MSM pass: {sort, accumulate, reduce, finish} (needs set_option("profile", 1))."""
        ms = (ctypes.c_float * 4)()
        rc = self.lib.kzg_gpu_phase_times(self.h, ms, 4)
        if rc != 0:
            return None
        return dict(zip(("sort", "accumulate", "reduce", "finish"), [float(v) for v in ms]))

    def msm_stats(self):
        out = (ctypes.c_uint64 * 6)()
        if self.lib.kzg_gpu_msm_stats(self.h, out, 6) != 0:
            return None
        return dict(zip(("records", "buckets", "tasks", "max_tasks_per_bucket", "window_bits", "task_chunk"), [int(v) for v in out]))

    def selftest(self):
        f = ctypes.c_int(-1)
        self._check(self.lib.kzg_gpu_selftest(self.h, ctypes.byref(f)), "selftest")
        return f.value


class Group:
    """kzg_gpu_group: several devices of one box behind the C ABI (one process).  Blob batches are sharded by blob,
    a single large MSM by point range (include/kzg_gpu.h "Multi-GPU")."""

    def __init__(self, curve, devices=None):
        self.lib = load_library()
        self.curve = CURVE_IDS[curve.lower()] if isinstance(curve, str) else int(curve)
        self.M = self.lib.kzg_gpu_modbytes(self.curve)
        self.SB = self.lib.kzg_gpu_scalar_bytes(self.curve)
        self.r = CURVE_ORDER[self.curve]
        h = ctypes.c_void_p()
        if devices is None or isinstance(devices, int):
            rc = self.lib.kzg_gpu_group_create(ctypes.byref(h), self.curve, None, int(devices or 0))
        else:
            arr = (ctypes.c_int * len(devices))(*devices)
            rc = self.lib.kzg_gpu_group_create(ctypes.byref(h), self.curve, arr, len(devices))
        if rc != 0:
            raise RuntimeError("kzg_gpu_group_create failed: %s (CUDA devices are required; no CPU fallback)"
                               % self.lib.kzg_gpu_strerror(rc).decode())
        self.h = h
        self.size = self.lib.kzg_gpu_group_size(h)

    def close(self):
        if getattr(self, "h", None):
            self.lib.kzg_gpu_group_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError("%s failed: %s [%s]" % (what, self.lib.kzg_gpu_strerror(rc).decode(),
                                                       self.lib.kzg_gpu_group_last_error(self.h).decode()))

    def set_option(self, name, value):
        self._check(self.lib.kzg_gpu_group_set_option(self.h, name.encode(), int(value)), "group set_option " + name)

    def srs_load(self, g1, g2=None):
        n = len(g1) // (2 * self.M)
        self._check(self.lib.kzg_gpu_group_srs_load(self.h, n, g1, g2), "group srs_load")

    def setup_powers(self, s, n, want_bytes=True):
        s_le = (int(s) % self.r).to_bytes(self.SB, "little")
        g1 = ctypes.create_string_buffer(n * 2 * self.M) if want_bytes else None
        g2 = ctypes.create_string_buffer(n * 4 * self.M) if want_bytes else None
        self._check(self.lib.kzg_gpu_group_setup_powers(self.h, s_le, n, g1, g2), "group setup_powers")
        return (g1.raw, g2.raw) if want_bytes else (None, None)

    def msm_g1(self, scalars, count=None, base_offset=0):
        count = len(scalars) // self.SB if count is None else count
        out = ctypes.create_string_buffer(2 * self.M)
        inf = ctypes.c_int(0)
        self._check(self.lib.kzg_gpu_group_msm_g1(self.h, _ptr(scalars), count, base_offset, out, ctypes.byref(inf)), "group msm_g1")
        return out.raw, bool(inf.value)

    def msm_g1_batch(self, scalars, count, batch):
        out = ctypes.create_string_buffer(batch * 2 * self.M)
        inf = (ctypes.c_int * batch)()
        self._check(self.lib.kzg_gpu_group_msm_g1_batch(self.h, _ptr(scalars), count, batch, out, inf), "group msm_g1_batch")
        return out.raw, [bool(v) for v in inf]

    def prove_batch(self, coeffs, n, first_roots, k):
        batch = len(first_roots)
        out = ctypes.create_string_buffer(batch * 2 * self.M)
        inf = (ctypes.c_int * batch)()
        roots = (ctypes.c_int64 * batch)(*first_roots)
        self._check(self.lib.kzg_gpu_group_prove_batch(self.h, _ptr(coeffs), n, roots, k, batch, out, inf), "group prove_batch")
        return out.raw, [bool(v) for v in inf]


# ------------------------------------------------------------------------------------------------
# kzg.h value types
class Poly:
    """kzg::poly (src/kzg.h:89-125): coefficient vector over Fr, low degree first, normalised like NTL's
    ZZ_pX (no trailing zeros).  Held as the flat little-endian SB-byte records the C ABI consumes."""

    def __init__(self, ctx_or_sizes, coeffs):
        self.SB, self.r = (ctx_or_sizes.SB, ctx_or_sizes.r) if hasattr(ctx_or_sizes, "SB") else ctx_or_sizes
        if isinstance(coeffs, (bytes, bytearray)):
            ints = [int.from_bytes(coeffs[i:i + self.SB], "little") % self.r for i in range(0, len(coeffs), self.SB)]
        else:
            ints = [int(c) % self.r for c in coeffs]
        while ints and ints[-1] == 0:
            ints.pop()
        self.coeffs = ints
        self.raw = b"".join(c.to_bytes(self.SB, "little") for c in ints)

    def __len__(self):
        return len(self.coeffs)

    @property
    def degree(self):
        return len(self.coeffs) - 1

    @classmethod
    def from_blob(cls, ctx, ys, offset=0):
        """poly::from_blob (src/poly.cpp:4-6 -> polyfit, src/util.cpp:180-184) for a blob whose chunk i sits at
        x = offset + i (blob::from_bytes / from_string): interpolation on the device."""
        raw = ys if isinstance(ys, (bytes, bytearray)) else b"".join((int(v) % ctx.r).to_bytes(ctx.SB, "little") for v in ys)
        coeffs, _ = ctx.interpolate_range(bytes(raw), offset, want_vanishing=False)
        return cls(ctx, coeffs)

    def serialize(self):
        """serialize_ZZ_pX (src/util.cpp:118-144): int64 degree | per coefficient [u8 nbytes | LE magnitude]."""
        out = [struct.pack("<q", self.degree)]
        for c in self.coeffs:
            nb = (c.bit_length() + 7) // 8
            out.append(bytes([nb]) + c.to_bytes(nb, "little"))
        return b"".join(out)

    @classmethod
    def deserialize(cls, sizes, raw):
        """deserialize_ZZ_pX (src/util.cpp:146-170)."""
        (deg,) = struct.unpack_from("<q", raw, 0)
        off, coeffs = 8, []
        for _ in range(deg + 1):
            nb = raw[off]
            coeffs.append(int.from_bytes(raw[off + 1:off + 1 + nb], "little"))
            off += 1 + nb
        return cls(sizes, coeffs)


class _CurvePoint:
    """kzg::commit / kzg::proof (src/kzg.h:127-180): one G1 point; serialize() = serialize_ECP
    (src/util.cpp:78-94): uint32 len | 0x04 | x | y, big-endian, infinity = all-zero coordinates."""

    def __init__(self, xy, is_inf):
        self.xy = bytes(xy)
        self.is_inf = bool(is_inf)

    def serialize(self):
        body = b"\x04" + self.xy
        return struct.pack("<I", len(body)) + body

    @classmethod
    def deserialize(cls, raw):
        (ln,) = struct.unpack_from("<I", raw, 0)
        xy = bytes(raw[5:4 + ln])
        return cls(xy, xy == bytes(len(xy)))

    def __eq__(self, other):
        return isinstance(other, _CurvePoint) and self.xy == other.xy


class Commit(_CurvePoint):
    pass


class Proof(_CurvePoint):
    pass


class TrustedSetup:
    """kzg::trusted_setup (src/kzg.h:182-290) with its _G1/_G2 vectors resident on the GPU."""

    def __init__(self, ctx, g1_bytes, g2_bytes):
        self.ctx = ctx
        self.g1_bytes, self.g2_bytes = g1_bytes, g2_bytes
        self.n = len(g1_bytes) // (2 * ctx.M)

    @classmethod
    def generate(cls, curve, num_coeff, secret, device=0, ctx=None):
        """trusted_setup(int num_coeff) (src/trusted_setup.cpp:21-74) with the secret supplied by the caller
        (the reference draws it from a CSPRNG and discards it, util.cpp:62-76)."""
        if num_coeff < 2:
            raise ValueError("num_coeff must be at least 2")
        ctx = ctx or Context(curve, device)
        g1, g2 = ctx.setup_powers(secret, num_coeff)
        return cls(ctx, g1, g2)

    @classmethod
    def from_file_bytes(cls, curve, raw, device=0, ctx=None):
        """trusted_setup(const std::string& filename) (src/trusted_setup.cpp:76-121): uint64 n, then n framed
        G1 octets and n framed G2 octets."""
        ctx = ctx or Context(curve, device)
        M = ctx.M
        if len(raw) < 8:
            raise RuntimeError("could not read number of coefficients from file")
        (n,) = struct.unpack_from("<Q", raw, 0)
        l1, l2 = 4 + 2 * M + 1, 4 + 4 * M + 1
        if len(raw) < 8 + n * (l1 + l2):
            raise RuntimeError("setup file is truncated")
        g1 = bytearray()
        off = 8
        for _ in range(n):
            g1 += raw[off + 5:off + l1]
            off += l1
        g2 = bytearray()
        for _ in range(n):
            g2 += raw[off + 5:off + l2]
            off += l2
        ctx.srs_load(bytes(g1), bytes(g2))
        return cls(ctx, bytes(g1), bytes(g2))

    def export_setup(self):
        """export_setup (src/trusted_setup.cpp:256-287)."""
        M = self.ctx.M
        out = [struct.pack("<Q", self.n)]
        for i in range(self.n):
            out.append(struct.pack("<I", 2 * M + 1) + b"\x04" + self.g1_bytes[i * 2 * M:(i + 1) * 2 * M])
        for i in range(self.n):
            out.append(struct.pack("<I", 4 * M + 1) + b"\x04" + self.g2_bytes[i * 4 * M:(i + 1) * 4 * M])
        return b"".join(out)

    def _poly(self, p):
        return p if isinstance(p, Poly) else Poly(self.ctx, p)

    def create_commit(self, poly):
        poly = self._poly(poly)
        if len(poly) >= self.n:  # deg(P)+1 < _G1.size()  (src/trusted_setup.cpp:138-139)
            raise ValueError("polynomial degree be at most one less than the setup size (num_coeffs)")
        xy, inf = self.ctx.msm_g1(poly.raw, len(poly), 0)
        return Commit(xy, inf)

    def verify_commit(self, commit, poly):
        return self.create_commit(poly) == commit

    def commit_bytes(self, data, chunk_size=None, byte_offset=0):
        """kzg-cli's commit_file path (demo/shared/kzg-cli.cpp:41-63) in one device pass: blob::from_bytes(data,
        byte_offset, len(data), chunk_size) -> poly::from_blob -> create_commit.  Returns (Commit, Poly)."""
        chunk_size = self.ctx.SB - 1 if chunk_size is None else chunk_size
        if chunk_size > self.ctx.SB - 1:  # src/blob.cpp:20-25
            raise ValueError("chunk_size must be at most MAX_CHUNK_BYTES.")
        if chunk_size < 1 or byte_offset % chunk_size != 0:
            raise ValueError("byte_offset is not a multiple of chunk_size.")
        if len(data) % chunk_size != 0:
            raise ValueError("byte_length is not a multiple of chunk_size.")
        if len(data) // chunk_size >= self.n:
            raise ValueError("polynomial degree be at most one less than the setup size (num_coeffs)")
        xy, inf, coeffs = self.ctx.commit_bytes(bytes(data), chunk_size, byte_offset // chunk_size)
        return Commit(xy, inf), Poly(self.ctx, coeffs or b"")

    def create_proof(self, poly, chunk_offset, chunk_length):
        poly = self._poly(poly)
        if chunk_length < 1:  # src/trusted_setup.cpp:215-216
            raise ValueError("chunk_length must be 1 or greater")
        if len(poly) - chunk_length >= self.n:
            raise ValueError("polynomial degree be at most one less than the setup size (num_coeffs)")
        xy, inf = self.ctx.prove(poly.raw, chunk_offset, chunk_length)
        return Proof(xy, inf)

    def create_proof_bytes(self, poly, byte_offset, byte_length, chunk_size):
        """create_proof(poly, byte_offset, byte_length, chunk_size) (src/trusted_setup.cpp:203-212)."""
        if chunk_size > self.ctx.SB - 1:
            raise ValueError("chunk_size must be at most MAX_CHUNK_BYTES.")
        if byte_offset % chunk_size:
            raise ValueError("byte_offset is not a multiple of chunk_size.")
        if byte_length % chunk_size:
            raise ValueError("byte_length is not a multiple of chunk_size.")
        return self.create_proof(poly, byte_offset // chunk_size, byte_length // chunk_size)

    def evaluate(self, poly, chunk_offset, chunk_length):
        """evaluate_polynomial_points (src/util.cpp:186-211): the opened values y_i = P(chunk_offset + i)."""
        poly = self._poly(poly)
        return self.ctx.poly_eval_range(poly.raw, chunk_offset, chunk_length)

    def verify_proof_msms(self, vanishing_coeffs, interp_coeffs):
        """The two MSMs of verify_proof (src/trusted_setup.cpp:241, 246): polyeval_G2(Z) and polyeval_G1(I).
        Z and I come from the host's linear_roots_and_polyfit (src/util.cpp:172-178); the pairing check that
        consumes the two points stays in the reference's host code (:242-253)."""
        Z, I = self._poly(vanishing_coeffs), self._poly(interp_coeffs)
        if len(Z) - 1 < 1:
            raise ValueError("expected_data.size() must be 1 or greater")
        z_xy, z_inf = self.ctx.msm_g2(Z.raw, len(Z), 0)
        i_xy, i_inf = self.ctx.msm_g1(I.raw, len(I), 0) if len(I) else (bytes(2 * self.ctx.M), True)
        return (z_xy, z_inf), (i_xy, i_inf)
