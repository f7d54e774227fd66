#!/usr/bin/env python
"""bench.py — headline benchmark of the KZG hot path on B200 (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--log2n 20]

Own arm (--impl b200, default).  One "step" = one BLS12-381 G1 Pippenger MSM over 2^log2n synthetic SRS
points per GPU (BASELINE.json config "G1 MSM 2^16-2^22 synthetic SRS", the size the metric is quoted on:
2^20).  SRS = [s^i]G1 generated on the device (seeded s); scalars uniform below r.
  value      points/s with scalars already resident in HBM: kzg_gpu_msm_g1_launch, CUDA events on the
             context's stream, L2 flushed (256 MiB memset) before every timed step.
  e2e        the same MSM through the host-buffer C-ABI call kzg_gpu_msm_g1 (what polyeval_G1 calls):
             pinned host scalars -> H2D -> MSM -> result bytes D2H, inside the timed region.
  N > 1      one process per GPU (torchrun), ONE SRS of N*2^log2n points: rank g holds the DISTINCT slice
             [g*2^log2n, (g+1)*2^log2n) (kzg_gpu_setup_powers_range) and its own scalars, computes its partial sum,
             and the N projective partials (4*L words each) are exchanged with one NCCL all-gather and folded on
             every rank.  Weak scaling; value = N*2^log2n / max-over-ranks time.
  parity     after the timed loops the folded result is ASSERTED equal to [P(s)]G1, P(s) = sum_g sum_i c_gi s^(g n + i)
             evaluated in Fr on the host from every rank's scalars (one scalar multiplication by the CPU oracle,
             used as the checker only, outside every timed region).  The run fails otherwise.
  strong     the configurations SURVEY.md 8(d) names as STRONG scaling: ONE 2^20-point and ONE 2^22-point MSM split N
             ways by contiguous point range, and 4096 blob commitments IN TOTAL sharded N ways (512 per GPU at N = 8),
             each with its own parity assertion.
  commit     batched blob commitments + single-point proofs (config 3) with roofline and CPU baseline objects.
             commit.direct_table: the same batch (and one single commitment) through the opt-in direct-lookup table
             (csrc/msm_direct.cuh, --direct-table-gb G, default 40 GiB of HBM; 0 skips it), outputs asserted byte-identical.
  secondary  the other kernels of the path: G2 MSM of 4097 terms (BN254, BLS12-381), quotient and proof at
             n = 8192 / k = 4096 (config 5), trusted setup of 2^20 indices.
Reference arm (--impl reference): the UNMODIFIED reference library (oracle/_ref/kzgref-bls12381, built from
/root/reference by oracle/build_ref.sh) runs trusted_setup::create_commit == polyeval_G1 on the host cores:
one process per host thread (the reference itself is single-threaded on this path), each a bounded sample
of the same workload (4096 of the points).  Rank 0 only.

The oracle (oracle/) is used here ONLY for the cpu_baseline / reference legs and as the post-run parity checker,
never on the product path and never inside a timed region.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))

CURVE = "bls12381"
FP_LIMBS = 12
SCALAR_BITS = 255
SECRET = 0x4B5A47_0123456789ABCDEF_FEDCBA9876543210_4B5A47  # seed 0x4b5a47
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "kzgref-" + CURVE)
TOP_LIMB = 0x73EDA753299D7D48  # top 64 bits of r: scalars are drawn uniformly below TOP_LIMB << 192 (< r)


# ------------------------------------------------------------------------------------------------
def canonical_counts(n, limbs=FP_LIMBS, bits=SCALAR_BITS):
    """canonical Pippenger modmul and IMAD count per point."""
    lg = n.bit_length() - 1
    c = min(16, max(8, lg - 4))
    W = -(-(bits + 1) // c)
    modmul = W * 10 + W * (1 << (c - 1)) * 2 * 14 / n
    imad = modmul * 2 * (2 * limbs * limbs + limbs)
    return {"c": c, "W": W, "modmul_per_point": modmul, "imad_per_point": imad}


def workload_config(args):
    """Same dict on both arms (the driver compares them)."""
    return {"workload": "BLS12-381 G1 MSM, 2^%d synthetic SRS points per GPU (polyeval_G1); N>1: point-range split + "
                        "NCCL all-gather of the partial sums + fold" % args.log2n,
            "curve": CURVE, "points_per_gpu": 1 << args.log2n, "canonical_window_bits": canonical_counts(1 << args.log2n)["c"],
            "l2": "256 MiB memset flush before every timed step; inputs 128 MiB"}


def random_scalars(n, seed):
    import numpy as np
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 1 << 63, size=(n, 4), dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=(n, 4), dtype=np.uint64)
    a[:, 3] = rng.integers(0, TOP_LIMB, size=n, dtype=np.uint64)
    return a  # little-endian 32-byte records when viewed as uint8


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._pump, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None
        self.t_lo = self.t_hi = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def mark_start(self):
        self.t_lo = time.time()

    def mark_stop(self):
        self.t_hi = time.time()

    def summary(self):
        if self.proc:
            self.proc.terminate()
        rows = [r for t, r in self.rows if self.t_lo is not None and self.t_lo - 0.05 <= t <= (self.t_hi or 1e30) + 0.15]
        rows = rows or [r for _, r in self.rows]
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": len(rows)}
        try:
            sm = [float(r[0]) for r in rows]
            out["sm_mhz"] = statistics.median(sm)
            out["sm_max_mhz"] = max(float(r[1]) for r in rows)
            out["power_w_max"] = max(float(r[2]) for r in rows)
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            out["reasons"] = [nm for i, nm in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        except Exception:
            pass
        return out


# ------------------------------------------------------------------------------------------------
def reference_ms(what, sample_points, procs, reps, workdir, srs_path, seed=1, extra=()):
    """Time the reference's create_commit / create_proof (src/trusted_setup.cpp:137-228) on `procs` host processes at
    once, each over a polynomial of `sample_points` coefficients.  Returns (ms per call of the slowest, wall s)."""
    import numpy as np
    paths = []
    for p in range(procs):
        path = os.path.join(workdir, "coeffs%d.bin" % p)
        random_scalars(sample_points, seed * 1000 + p).view(np.uint8).tofile(path)
        paths.append(path)
    t0 = time.time()
    ps = [subprocess.Popen([REF_BIN, "time", srs_path, path, what, str(reps), *map(str, extra)], stdout=subprocess.PIPE, text=True)
          for path in paths]
    ms = [float(p.communicate()[0].strip()) for p in ps]
    wall = time.time() - t0
    if any(p.returncode for p in ps):
        raise RuntimeError("reference harness failed")
    return max(ms), wall


def reference_points_per_s(sample_points, procs, reps, workdir, srs_path, seed=1):
    worst, wall = reference_ms("commit", sample_points, procs, reps, workdir, srs_path, seed)
    return procs * sample_points / (worst * 1e-3), worst, wall


def make_reference_srs(n, workdir):
    """An SRS file written by the reference's own trusted_setup(n) + export_setup (random secret; not timed)."""
    path = os.path.join(workdir, "srs.bin")
    subprocess.run([REF_BIN, "setup", str(n), path], check=True)
    return path


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if not os.path.exists(REF_BIN):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/kzgref-bls12381 not built (run oracle/build_ref.sh)"}))
        return 0
    cores = os.cpu_count() or 1
    sample = args.ref_sample
    with tempfile.TemporaryDirectory() as td:
        srs = make_reference_srs(sample + 1, td)
        for w in range(args.warmup):
            reference_points_per_s(256, cores, 1, td, srs, seed=90 + w)
        t_total, pts = 0.0, 0
        for k in range(args.steps):
            v, worst_ms, _ = reference_points_per_s(sample, cores, 1, td, srs, seed=k + 1)
            t_total += worst_ms
            pts += cores * sample
        # the other half of the metric: blob commitments and single-point proofs per second at 4096 coefficients
        commit_ms, _ = reference_ms("commit", 4096, cores, 1, td, srs, seed=501) if sample >= 4096 else (None, 0)
        prove_ms, _ = reference_ms("prove", 4096, cores, 1, td, srs, seed=502, extra=(17, 1)) if sample >= 4096 else (None, 0)
    value = pts / (t_total * 1e-3)
    line = {
        "impl": "reference", "metric": "g1_msm_points_per_s", "value": value, "unit": "points/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32 limbs (reference: 58-bit limbs in int64)", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": "points/s", "cores": cores, "kind": "reference",
                         "sample": "%d host processes x create_commit over %d of the 2^%d terms per step (reference MSM is "
                                   "strictly per-term, so points/s is size-independent); harness-level parallelism, the "
                                   "reference itself is single-threaded here.  ms_per_step is the time of ONE SAMPLE "
                                   "(%d terms per process), not of a 2^%d-term MSM"
                                   % (cores, sample, args.log2n, sample, args.log2n)},
        "e2e": {"value": value, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if commit_ms:
        line["commit"] = {"metric": "blob_commits_per_s", "unit": "commits/s", "coeffs_per_blob": 4096, "cores": cores,
                          "value": cores / (commit_ms * 1e-3), "e2e": cores / (commit_ms * 1e-3), "ms_per_commit_one_core": commit_ms,
                          "proofs_per_s_e2e": cores / (prove_ms * 1e-3), "ms_per_proof_one_core": prove_ms,
                          "sample": "one create_commit and one create_proof(poly, 17, 1) of a 4096-coefficient polynomial per "
                                    "host process, %d processes at once" % cores}
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
def scalars_to_ints(arr):
    """(n, 4) uint64 little-endian limbs -> list of Python ints."""
    raw = arr.tobytes()
    return [int.from_bytes(raw[i:i + 32], "little") for i in range(0, len(raw), 32)]


def _timed(fn):
    """milliseconds of one host-timed call"""
    t0 = time.perf_counter()
    fn()
    return (time.perf_counter() - t0) * 1e3


def fr_eval(ints, s, r, first=0):
    """sum_i ints[i] * s^(first + i) mod r (Horner)."""
    acc = 0
    for c in reversed(ints):
        acc = (acc * s + c) % r
    return acc * pow(s, first, r) % r


def column_sums(arr, nblobs, ncoef):
    """sum over blobs of coefficient i, for every i: arr is (nblobs*ncoef, 4) uint64 -> ncoef Python ints."""
    import numpy as np
    a32 = arr.view(np.uint32).reshape(nblobs, ncoef, 8).astype(np.uint64).sum(axis=0)  # < 2^32 * nblobs per limb
    return [sum(int(a32[i, j]) << (32 * j) for j in range(8)) for i in range(ncoef)]


def run_b200(args):
    import numpy as np
    import torch
    import kzg_b200
    import sharding

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the KZG hot path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    # Keep stdout to the one JSON line: libraries (NCCL prints its version banner to fd 1) write to stderr instead.
    sys.stdout.flush()
    json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather_bytes(b):
        """every rank's fixed-size byte string, in rank order"""
        if dist is None:
            return [bytes(b)]
        t = torch.frombuffer(bytearray(b), dtype=torch.uint8).to(dev)
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [bytes(o.cpu().numpy().tobytes()) for o in out]

    ctx = kzg_b200.Context(CURVE, local)
    R_ORDER = ctx.r
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    ctx.set_option("setup_g2", 0)
    ctx.set_option("profile", 1)
    ctx.set_option("precompute", args.precompute)
    ctx.set_option("window_bits", args.window_bits)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def oracle_point(e):
        """[e]G1 as the 2M affine bytes the C ABI returns — the CPU oracle as the CHECKER (never timed, never on the product path)"""
        sys.path.insert(0, ROOT)
        from oracle.kzg_oracle import get_curve
        C = get_curve(CURVE)
        return C.g1_to_octet(C.mul(e % C.r, C.g1))[1:]

    def split_msm(n_local, first_index, seed, steps, label, sampler=None):
        """One MSM whose points [first_index, first_index + n_local) and scalars live on this rank: resident and
        end-to-end timings (max over ranks), launch/phase statistics, and the parity assertion."""
        ctx.set_option("finish_affine", 0 if world > 1 else 1)
        ctx.setup_powers(SECRET, n_local, want_bytes=False, first_index=first_index)
        h_np = random_scalars(n_local, seed)
        h_scalars = torch.from_numpy(h_np.view(np.uint8).reshape(-1)).pin_memory()
        d_scalars = h_scalars.cuda(non_blocking=False)

        def fold_partials():
            return sharding.gather_partials_device(dist, ctx, dev)

        def step_resident():
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            flush.zero_()
            torch.cuda.synchronize()
            e0.record(stream)
            ctx.msm_g1_launch(d_scalars, n_local, 1, 0)
            if world > 1:  # MSM -> NCCL all-gather -> fold, all ordered on the context's stream and inside the events
                with torch.cuda.stream(stream):
                    fold_partials()  # (the per-rank MSM skipped its own affine normalisation: finish_affine = 0)
            e1.record(stream)
            e1.synchronize()
            return e0.elapsed_time(e1)

        def step_e2e():
            flush.zero_()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            # the call polyeval_G1 makes in the drop-in: host scalars in (chunked H2D under the digit kernel), result out;
            # N > 1: the same call leaves the projective partial, then NCCL all-gather + fold + normalise + bytes to host
            xy, inf = ctx.msm_g1(h_scalars, n_local, 0)
            if world > 1:
                with torch.cuda.stream(stream):
                    xy, inf = fold_partials()
            return (time.perf_counter() - t0) * 1e3, xy

        for _ in range(max(args.warmup, 3)):
            step_resident()
        barrier()
        if sampler:
            sampler.mark_start()
        l_before = ctx.launch_count
        step_ms, phase_log = [], []
        for _ in range(steps):
            step_ms.append(step_resident())
            phase_log.append(ctx.phase_times())
        barrier()
        launches = ctx.launch_count - l_before
        stats = ctx.msm_stats()
        xy_res = None
        if world == 1:
            xy_res, _ = ctx.msm_fetch(1)
        e2e_ms, xy = [], None
        step_e2e()
        for _ in range(steps):
            ms, xy = step_e2e()
            e2e_ms.append(ms)
        barrier()
        if sampler:
            sampler.mark_stop()
        # ---- parity (outside every timed region): folded bytes == [sum_g sum_i c_gi s^(first_g + i)] G1
        e_mine = fr_eval(scalars_to_ints(h_np), SECRET % R_ORDER, R_ORDER, first_index)
        parts = gather_bytes(e_mine.to_bytes(32, "little"))
        got = gather_bytes(xy)
        if rank == 0:
            want = oracle_point(sum(int.from_bytes(p, "little") for p in parts))
            assert all(g == want for g in got), "%s: folded MSM result differs from [P(s)]G1" % label
            assert xy_res is None or xy_res == want, "%s: resident result differs from [P(s)]G1" % label
        phases = {k: sum(p[k] for p in phase_log) / len(phase_log) for k in phase_log[0]} if phase_log and phase_log[0] else {}
        del d_scalars
        return {"ms": allmax(sum(step_ms)) / steps, "e2e_ms": allmax(sum(e2e_ms)) / steps, "launches": launches, "stats": stats,
                "phases": phases, "parity_asserted": True}

    # ---------------- headline: 2^log2n points PER GPU (weak scaling), rank g = slice g of one N * 2^log2n-point SRS
    n = 1 << args.log2n
    sampler = ClockSampler(local) if rank == 0 else None
    head = split_msm(n, rank * n, 0x4B5A47 + rank, args.steps, "headline", sampler)
    ms_per_step, e2e_ms_per_step = head["ms"], head["e2e_ms"]
    phases, stats, gpu_launches = head["phases"], head["stats"], head["launches"]

    # ---------------- strong scaling: ONE MSM split N ways, 4096 blobs IN TOTAL
    strong = {}
    if not args.no_strong:
        for lg in [int(x) for x in args.strong_log2n.split(",") if x]:
            if lg == args.log2n and world == 1:
                strong["msm_2p%d" % lg] = {"points": n, "ms": ms_per_step, "points_per_s": n / (ms_per_step * 1e-3),
                                           "e2e_ms": e2e_ms_per_step, "e2e_points_per_s": n / (e2e_ms_per_step * 1e-3),
                                           "parity_asserted": True, "note": "same measurement as the headline at N = 1"}
                continue
            tot = 1 << lg
            lo, hi = sharding.point_range(rank, world, tot)
            r = split_msm(hi - lo, lo, 0x5A0000 + 64 * lg + rank, max(3, min(args.steps, 5)), "strong 2^%d" % lg)
            strong["msm_2p%d" % lg] = {"points": tot, "points_per_gpu": tot // world, "ms": r["ms"], "points_per_s": tot / (r["ms"] * 1e-3),
                                       "e2e_ms": r["e2e_ms"], "e2e_points_per_s": tot / (r["e2e_ms"] * 1e-3),
                                       "window_bits": r["stats"]["window_bits"] if r["stats"] else None,
                                       "phases_ms": r["phases"], "parity_asserted": True}

    # ---------------- batched blob commitments + proofs (config 3): args.blobs blobs IN TOTAL, sharded by blob, no collective
    commit = None
    if args.blobs > 0:
        nc = 4096
        b0, b1 = sharding.blob_slice(rank, world, args.blobs)
        nb = b1 - b0
        ctx.set_option("window_bits", 0)
        ctx.set_option("finish_affine", 1)
        ctx.setup_powers(SECRET, nc + 1, want_bytes=False)
        hb_np = random_scalars(max(nb, 1) * nc, 77 + rank)
        hb = torch.from_numpy(hb_np.view(np.uint8).reshape(-1)).pin_memory()
        db = hb.cuda()
        times = []
        cstats = None
        if nb:
            for it in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                flush.zero_()
                torch.cuda.synchronize()
                e0.record(stream)
                ctx.msm_g1_launch(db, nc, nb, 0)
                e1.record(stream)
                e1.synchronize()
                times.append(e0.elapsed_time(e1))
            cstats = ctx.msm_stats()
            cphases = ctx.phase_times()
            ctx.msm_g1_batch(hb, nc, nb)  # untimed: grows the staging buffers for this batch size
            barrier()
            t0 = time.perf_counter()
            out_xy, out_inf = ctx.msm_g1_batch(hb, nc, nb)
            t_e2e = (time.perf_counter() - t0) * 1e3
            # batched single-point proofs (create_proof(poly, off, 1) for every blob): quotient + MSM on device
            roots = [int(i % 4096) for i in range(nb)]
            ctx.prove_batch(hb, nc, roots, 1)
            t0 = time.perf_counter()
            ctx.prove_batch(hb, nc, roots, 1)
            t_prove = (time.perf_counter() - t0) * 1e3
        else:
            times, t_e2e, t_prove, out_xy, cphases = [0.0, 0.0, 0.0], 0.0, 0.0, b"", None
            barrier()
        # parity: checksum of checksums  sum_b commit_b == [sum_b P_b(s)] G1  (host additions by the oracle's group law)
        sys.path.insert(0, ROOT)
        from oracle.kzg_oracle import get_curve
        C = get_curve(CURVE)
        acc = None
        for b in range(nb):
            body = out_xy[b * 96:(b + 1) * 96]
            pt = None if body == bytes(96) else (int.from_bytes(body[:48], "big"), int.from_bytes(body[48:], "big"))
            acc = C.add(acc, pt)
        e_mine = fr_eval(column_sums(hb_np, nb, nc), SECRET % R_ORDER, R_ORDER) if nb else 0
        sums = gather_bytes(C.g1_to_octet(acc)[1:])
        evals = gather_bytes(e_mine.to_bytes(40, "little"))
        if rank == 0:
            tot_pt = None
            for sb_ in sums:
                pt = None if sb_ == bytes(96) else (int.from_bytes(sb_[:48], "big"), int.from_bytes(sb_[48:], "big"))
                tot_pt = C.add(tot_pt, pt)
            want = C.mul(sum(int.from_bytes(e, "little") for e in evals) % C.r, C.g1)
            assert tot_pt == want, "batched commitments: sum of all commitments differs from [sum_b P_b(s)]G1"
        commit = {"blobs": args.blobs, "blobs_per_gpu": nb, "coeffs": nc, "ms_resident": allmax(min(times[1:])),
                  "ms_e2e": allmax(t_e2e), "ms_prove_e2e": allmax(t_prove), "stats": cstats, "phases": cphases}
        # ---- the same batch through the DIRECT-LOOKUP table (csrc/msm_direct.cuh; opt-in: tens of GiB of HBM per SRS):
        # every multiple a digit can select is tabulated, so a commitment is a plain sum of looked-up points.  Outputs
        # must be byte-identical to the bucket pipeline's, whose checksum of checksums was asserted above.
        if args.direct_table_gb > 0:
            dt, d_e2e, d_prove, d_single, b_single, t_build, dstats = [0.0, 0.0, 0.0], 0.0, 0.0, 0.0, 0.0, 0.0, None
            if nb:
                one = hb[:nc * 32]
                ctx.msm_g1(one, nc)
                b_single = min(_timed(lambda: ctx.msm_g1(one, nc)) for _ in range(7))
                ctx.set_option("direct_table_gb", args.direct_table_gb)
                t0 = time.perf_counter()
                ctx.setup_powers(SECRET, nc + 1, want_bytes=False)
                t_build = (time.perf_counter() - t0) * 1e3
                dt = []
                for it in range(3):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    flush.zero_()
                    torch.cuda.synchronize()
                    e0.record(stream)
                    ctx.msm_g1_launch(db, nc, nb, 0)
                    e1.record(stream)
                    e1.synchronize()
                    dt.append(e0.elapsed_time(e1))
                dstats = ctx.msm_stats()
                ctx.msm_g1_batch(hb, nc, nb)
                t0 = time.perf_counter()
                d_xy, d_inf = ctx.msm_g1_batch(hb, nc, nb)
                d_e2e = (time.perf_counter() - t0) * 1e3
                assert d_xy == out_xy and list(d_inf) == list(out_inf), "direct-lookup table: commitments differ from the bucket pipeline's"
                ctx.prove_batch(hb, nc, roots, 1)
                t0 = time.perf_counter()
                ctx.prove_batch(hb, nc, roots, 1)
                d_prove = (time.perf_counter() - t0) * 1e3
                ctx.msm_g1(one, nc)
                d_single = min(_timed(lambda: ctx.msm_g1(one, nc)) for _ in range(7))
                ctx.set_option("direct_table_gb", 0)
            commit["direct"] = {"table_gib_budget": args.direct_table_gb, "window_bits": dstats["window_bits"] if dstats else None,
                                "table_build_ms": t_build, "ms_resident": allmax(min(dt[1:])), "ms_e2e": allmax(d_e2e),
                                "ms_prove_e2e": allmax(d_prove), "ms_one_commit_e2e": d_single, "ms_one_commit_e2e_buckets": b_single}
        del db

    # ---------------- secondary kernels (rank 0, N = 1 only): G2 MSM, quotient / proof at k = 4096, trusted setup
    secondary = None
    if world == 1 and not args.no_secondary:
        secondary = run_secondary(kzg_b200, np, torch, local)

    if rank == 0:
        points = n * world
        value = points / (ms_per_step * 1e-3)
        e2e_value = points / (e2e_ms_per_step * 1e-3)
        cc = canonical_counts(n)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        clocks = sampler.summary()
        sm_max = clocks.get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
        # Peak of the integer-multiply pipe: 16 lanes/clk/SMSP x 4 SMSP x 148 SMs IMAD-class issue slots/s; a 32x32->64
        # product (IMAD.WIDE) takes two slots.  The Fp modmul microbenchmark (600 slots per modmul, registers only)
        # is run beside it to show how much of that nominal figure dependent carry chains can reach.
        nominal = 64 * 148 * sm_max * 1e6
        wide = ctx.microbench(0)          # independent mad.lo.cc/madc.hi.cc pairs: 32x32->64 products/s
        modmul_rate = ctx.microbench(2)   # Fp Montgomery products/s
        imad_per_modmul = 2 * (2 * FP_LIMBS * FP_LIMBS + FP_LIMBS)
        measured = max(2.0 * wide, modmul_rate * imad_per_modmul)
        peak = nominal
        # dominant kernel: msm_accumulate, one launch per step; algorithmic work = one mixed XYZZ addition
        # (10 modmuls, ec.cuh madd) per record, records = points x windows
        acc_ms = phases["accumulate"]
        achieved = stats["records"] * 10 * imad_per_modmul / (acc_ms * 1e-3)
        step_canonical = value / world * cc["imad_per_point"]
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        hbm_ach = value / world * 128 / 1e9
        traffic, traffic_src = None, None
        try:  # dram bytes of one msm_accumulate launch: only from a committed ncu capture of this exact workload
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["msm_accumulate"]
            if tr["log2n"] == args.log2n and tr["window_bits"] == stats["window_bits"] and tr["records"] == stats["records"]:
                traffic, traffic_src = tr["dram_bytes_per_launch"], tr["source"]
        except Exception:
            pass
        line = {
            "metric": "g1_msm_points_per_s", "value": value, "unit": "points/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32 limbs (Montgomery, 12x32-bit Fp / 8x32-bit Fr)", "data": "synthetic",
            "config": workload_config(args),
            "e2e": {"value": e2e_value, "unit": "points/s", "h2d_bytes_per_step": n * 32, "d2h_bytes_per_step": 96 + 4,
                    "ms_per_step": e2e_ms_per_step},
            "gpu_launches": gpu_launches, "launches_per_msm": gpu_launches / args.steps,
            "parity_asserted": True,
            "parity_check": "folded result == [sum_g sum_i c_gi s^(g n + i)] G1 (Fr evaluation on the host + one oracle scalar "
                            "multiplication, after the timed region)",
            "roofline": {"bound": "int32-imad", "kernel": "msm_accumulate", "achieved": achieved / 1e12, "peak": peak / 1e12,
                         "unit": "TIMAD/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "traffic_unit": "bytes/launch", "algorithmic_bytes_per_launch": stats["records"] * 96 + stats["records"] * 4,
                         "kernel_ms": acc_ms, "kernel_share_of_step": acc_ms / ms_per_step,
                         "records_per_launch": stats["records"], "imad_per_record": 10 * imad_per_modmul,
                         "peak_source": "nominal: 64 IMAD-class slots/clk/SM x 148 SMs x sm_max_mhz (IMAD.WIDE = 2 slots)",
                         "peak_microbench": measured / 1e12,
                         "peak_microbench_source": "max(2 x IMAD.WIDE products/s, 600 x Fp modmuls/s) measured in this run",
                         "step_canonical": {"achieved": step_canonical / 1e12, "frac": step_canonical / peak,
                                            "imad_per_point": cc["imad_per_point"], "modmul_per_point": cc["modmul_per_point"],
                                            "note": "whole step against SURVEY 8(d)'s canonical Pippenger count (c=16, no "
                                                    "precomputed table); the window table does fewer operations than that"},
                         "hbm": {"achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_ach / hbm_peak,
                                 "bytes_per_point": 128, "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}},
            "clocks": clocks,
        }
        if phases:
            line["phases_ms"] = phases
        if strong:
            line["strong"] = strong
        if commit:
            nbt, cms = commit["blobs"], commit["ms_resident"]
            c4 = canonical_counts(commit["coeffs"])
            canon_imad = c4["imad_per_point"] * commit["coeffs"]          # 0.855 G IMAD per commitment
            cs = commit["stats"] or {}
            nbk = 1 << ((cs.get("window_bits") or 1) - 1)
            exec_imad = (cs.get("records", 0) * 10 + commit["blobs_per_gpu"] * 2 * nbk * 14) * imad_per_modmul  # rank 0's share
            line["commit"] = {"metric": "blob_commits_per_s", "value": nbt / (cms * 1e-3),
                              "e2e": nbt / (commit["ms_e2e"] * 1e-3), "unit": "commits/s",
                              "blobs": nbt, "blobs_per_gpu": commit["blobs_per_gpu"], "coeffs_per_blob": commit["coeffs"],
                              "scaling": "strong (%d blobs in total, sharded by blob, no collective)" % nbt,
                              "ms_resident": cms, "ms_e2e": commit["ms_e2e"],
                              "proofs_per_s_e2e": nbt / (commit["ms_prove_e2e"] * 1e-3),
                              "ms_prove_e2e": commit["ms_prove_e2e"], "parity_asserted": True,
                              "parity_check": "sum of all commitments == [sum_b P_b(s)] G1 (checksum of checksums)",
                              "phases_ms": commit["phases"],
                              "direct_table": None if "direct" not in commit else dict(
                                  commit["direct"], commits_per_s=nbt / (commit["direct"]["ms_resident"] * 1e-3),
                                  commits_per_s_e2e=nbt / (commit["direct"]["ms_e2e"] * 1e-3),
                                  proofs_per_s_e2e=nbt / (commit["direct"]["ms_prove_e2e"] * 1e-3), bytes_identical_to_bucket_pipeline=True,
                                  note="opt-in (kzg_gpu_set_option direct_table_gb): every multiple d*2^(cw)*G_i tabulated in HBM; no sort, "
                                       "no buckets, no weighted reduction"),
                              "roofline": {"bound": "int32-imad", "peak": peak / 1e12, "unit": "TIMAD/s",
                                           "executed": {"imad_per_gpu": exec_imad, "achieved": exec_imad / (cms * 1e-3) / 1e12,
                                                        "frac": exec_imad / (cms * 1e-3) / peak,
                                                        "note": "records x 10 modmul (mixed additions) + 2 x buckets x 14 modmul "
                                                                "(bucket reduction) per blob, x 600 IMAD, rank 0's share / its time"},
                                           "canonical": {"imad_per_commit": canon_imad,
                                                         "achieved": canon_imad * nbt / world / (cms * 1e-3) / 1e12,
                                                         "frac": canon_imad * nbt / world / (cms * 1e-3) / peak,
                                                         "commits_per_s_at_peak_per_gpu": peak / canon_imad}}}
        if secondary:
            for v in secondary.values():
                if isinstance(v, dict) and v.get("imad"):
                    v["frac"] = v["imad"] / (v["ms"] * 1e-3) / peak
            line["secondary"] = secondary
        # CPU baseline: the unmodified reference on the host cores, bounded sample (N = 1 only)
        if world == 1 and not args.no_cpu and os.path.exists(REF_BIN):
            try:
                with tempfile.TemporaryDirectory() as td:
                    sample = args.ref_sample
                    srs = make_reference_srs(max(sample, 4096) + 1, td)
                    v, worst_ms, wall = reference_points_per_s(sample, 1, 1, td, srs)
                    line["cpu_baseline"] = {"value": v, "unit": "points/s", "cores": 1, "kind": "reference",
                                            "sample": "reference create_commit (polyeval_G1) over %d of the 2^%d terms, 1 thread "
                                                      "(the reference MSM is single-threaded and strictly per-term): %.0f ms"
                                                      % (sample, args.log2n, worst_ms)}
                    if commit:
                        cm, _ = reference_ms("commit", 4096, 1, 1, td, srs, seed=501)
                        pm, _ = reference_ms("prove", 4096, 1, 1, td, srs, seed=502, extra=(17, 1))
                        line["commit"]["cpu_baseline"] = {"value": 1e3 / cm, "unit": "commits/s", "cores": 1, "kind": "reference",
                                                          "proofs_per_s": 1e3 / pm,
                                                          "sample": "one reference create_commit (%.0f ms) and one create_proof(poly, 17, 1) "
                                                                    "(%.0f ms) of a 4096-coefficient polynomial, 1 thread" % (cm, pm)}
            except Exception as e:  # the baseline leg must never sink the GPU numbers
                line["cpu_baseline"] = {"value": None, "unit": "points/s", "cores": 1, "kind": "reference", "sample": "failed: %r" % (e,)}
        json_out.write(json.dumps(line) + "\n")
        json_out.flush()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def run_secondary(kzg_b200, np, torch, local):
    """The other kernels of the path, each as {ms, modmul, imad}: the caller adds frac = imad / ms / peak."""
    out = {}

    def best_ms(fn, reps=5):
        fn()
        ts = []
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            ts.append((time.perf_counter() - t0) * 1e3)
        return min(ts)

    for cname, limbs, fr_limbs in (("bn254", 8, 8), ("bls12381", 12, 8)):
        c = kzg_b200.Context(cname, local)
        imad_fp = 2 * (2 * limbs * limbs + limbs)
        imad_fr = 2 * (2 * fr_limbs * fr_limbs + fr_limbs)
        c.setup_powers(SECRET, 8200 if cname == "bn254" else 4100, want_bytes=False)
        rng = np.random.default_rng(5 + limbs)
        sc = rng.integers(0, 256, size=(8192, c.SB), dtype=np.uint8)
        sc[:, c.SB - 1] &= 0x0F   # below r for both curves
        raw = sc.tobytes()
        # verify side: G2 MSM over the k + 1 = 4097 coefficients of Z (4097 (32 x 28) + 32 x 128 x 2 x 40 modmul)
        mm = 4097 * 32 * 28 + 32 * 128 * 2 * 40
        out["g2_msm_4097_" + cname] = {"ms": best_ms(lambda: c.msm_g2(raw[:4097 * c.SB], 4097)), "modmul": mm, "imad": mm * imad_fp,
                                       "what": "kzg_gpu_msm_g2, host scalars in, affine bytes out"}
        mm1 = 4097 * 32 * 10 + 32 * 128 * 2 * 14
        out["g1_msm_4097_" + cname] = {"ms": best_ms(lambda: c.msm_g1(raw[:4097 * c.SB], 4097)), "modmul": mm1, "imad": mm1 * imad_fp,
                                       "what": "kzg_gpu_msm_g1 (one create_commit of degree 4096), host scalars in, affine bytes out"}
        if cname == "bn254":  # config 5 as SURVEY 8(d) states it: degree 8191, k = 4096, q != 0
            mq = 4096 * 4096
            out["quotient_n8192_k4096_bn254"] = {"ms": best_ms(lambda: c.poly_quotient(raw[:8192 * c.SB], 0, 4096)), "modmul": mq,
                                                 "imad": mq * imad_fr, "what": "kzg_gpu_poly_quotient, canonical (n-k) k Fr modmuls"}
            out["prove_n8192_k4096_bn254"] = {"ms": best_ms(lambda: c.prove(raw[:8192 * c.SB], 0, 4096)), "modmul": mq + 4096 * 348,
                                              "imad": mq * imad_fr + 4096 * 348 * imad_fp,
                                              "what": "kzg_gpu_prove: quotient + G1 MSM of the 4096 quotient coefficients"}
        c.close()
    c = kzg_b200.Context("bls12381", local)
    c.set_option("precompute", 0)
    n = 1 << 20
    t = best_ms(lambda: c.setup_powers(SECRET, n, want_bytes=False), reps=2)
    out["setup_2p20_g1_g2_bls12381"] = {"ms": t, "modmul": n * 1216, "imad": n * 1216 * 600, "indices_per_s": n / (t * 1e-3),
                                        "what": "kzg_gpu_setup_powers, 2^20 indices, G1 and G2 (fixed-base comb; no MSM window table)"}
    c.close()
    return out


def run_single_process(args):
    """--single-process: ONE process drives --gpus N devices through the kzg_gpu_group C ABI (include/kzg_gpu.h
    "Multi-GPU": one context + one host worker thread per device) instead of one process per GPU over NCCL.  Prints one
    JSON line of the same shape; value = points/s of ONE 2^log2n-point MSM from pinned host memory, split by point
    range over the devices (strong scaling), plus the 4096-blob commit / proof batch sharded by blob.  Results are
    asserted byte-identical to a one-device group."""
    import numpy as np
    import torch
    import kzg_b200
    n, G = 1 << args.log2n, args.gpus
    big = torch.from_numpy(random_scalars(n, 0x4B5A47).view(np.uint8).reshape(-1)).pin_memory()
    nc, nb = 4096, max(args.blobs, 1)
    blobs = torch.from_numpy(random_scalars(nb * nc, 77).view(np.uint8).reshape(-1)).pin_memory()
    roots = [i % 4096 for i in range(nb)]

    def measure(g):
        grp = kzg_b200.Group(CURVE, g)
        grp.set_option("setup_g2", 0)
        grp.setup_powers(SECRET, n, want_bytes=False)
        for _ in range(max(args.warmup, 3)):
            grp.msm_g1(big, n)
        ts = []
        for _ in range(args.steps):
            t0 = time.perf_counter()
            xy = grp.msm_g1(big, n)
            ts.append((time.perf_counter() - t0) * 1e3)
        grp.setup_powers(SECRET, nc + 1, want_bytes=False)
        grp.msm_g1_batch(blobs, nc, nb)
        t0 = time.perf_counter()
        cm = grp.msm_g1_batch(blobs, nc, nb)
        t_c = (time.perf_counter() - t0) * 1e3
        grp.prove_batch(blobs, nc, roots, 1)
        t0 = time.perf_counter()
        pr = grp.prove_batch(blobs, nc, roots, 1)
        t_p = (time.perf_counter() - t0) * 1e3
        grp.close()
        return sum(ts) / len(ts), xy, t_c, cm, t_p, pr

    ms1, xy1, c1, cm1, p1, pr1 = measure(1)
    msG, xyG, cG, cmG, pG, prG = measure(G) if G > 1 else (ms1, xy1, c1, cm1, p1, pr1)
    assert xy1 == xyG and cm1 == cmG and pr1 == prG, "group results differ from the single-device results"
    line = {"metric": "g1_msm_points_per_s", "value": n / (msG * 1e-3), "unit": "points/s", "n_gpus": G, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": msG, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u32 limbs (Montgomery, 12x32-bit Fp / 8x32-bit Fr)", "data": "synthetic",
            "mode": "single process, kzg_gpu_group C ABI (one host worker thread per device)",
            "config": {"workload": "ONE BLS12-381 G1 MSM over 2^%d points from pinned host memory, split by point range over "
                                   "%d devices; partials folded on device 0" % (args.log2n, G), "curve": CURVE},
            "e2e": {"value": n / (msG * 1e-3), "unit": "points/s", "h2d_bytes_per_step": n * 32, "d2h_bytes_per_step": 96 + 4 + 192 * G,
                    "ms_per_step": msG},
            "one_device": {"ms_per_step": ms1, "speedup": ms1 / msG},
            "commit": {"blobs": nb, "coeffs_per_blob": nc, "e2e": nb / (cG * 1e-3), "unit": "commits/s", "ms_e2e": cG,
                       "one_device_ms": c1, "speedup": c1 / cG, "proofs_per_s_e2e": nb / (pG * 1e-3), "ms_prove_e2e": pG,
                       "prove_speedup": p1 / pG},
            "bytes_identical_to_one_device": True}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--log2n", type=int, default=20)
    ap.add_argument("--window-bits", type=int, default=0)
    ap.add_argument("--precompute", type=int, default=1)
    ap.add_argument("--blobs", type=int, default=4096,
                    help="blobs IN TOTAL for the batched-commit object, sharded over the GPUs (0 = skip); 4096 is BASELINE config 3")
    ap.add_argument("--ref-sample", type=int, default=16384, help="terms per reference process per step")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--strong-log2n", default="20,22", help="sizes of the single MSMs split N ways (strong scaling)")
    ap.add_argument("--no-secondary", action="store_true")
    ap.add_argument("--direct-table-gb", type=int, default=40,
                    help="GiB of HBM the direct-lookup table of the 4097-point commit SRS may take (0: skip that measurement)")
    ap.add_argument("--single-process", action="store_true",
                    help="drive --gpus N devices from ONE process through the kzg_gpu_group C ABI (no torchrun, no NCCL)")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.ref_sample == 16384:
            args.ref_sample = 4096
        return run_reference(args)
    if args.single_process:
        return run_single_process(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
