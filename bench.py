#!/usr/bin/env python
"""bench.py — headline benchmark of the KZG hot path on B200 (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--log2n 20]

Own arm (--impl b200, default).  One "step" = one BLS12-381 G1 Pippenger MSM over 2^log2n synthetic SRS
points per GPU (BASELINE.json config "G1 MSM 2^16-2^22 synthetic SRS", the size the metric is quoted on:
2^20).  SRS = [s^i]G1 generated on the device by kzg_gpu_setup_powers (seeded s); scalars uniform below r.
  value      points/s with scalars already resident in HBM: kzg_gpu_msm_g1_launch, CUDA events on the
             context's stream, L2 flushed (256 MiB memset) before every timed step.
  e2e        the same MSM through the host-buffer C-ABI call kzg_gpu_msm_g1 (what polyeval_G1 calls):
             pinned host scalars -> H2D -> MSM -> 2M result bytes D2H, inside the timed region.
  N > 1      one process per GPU (torchrun).  The path shards by POINT RANGE: rank g owns points
             [g*2^log2n, (g+1)*2^log2n) of one N*2^log2n-point MSM, computes its partial sum, and the N
             projective partials (4*L words each) are exchanged with one NCCL all-gather and folded on
             every rank (kzg_gpu_g1_sum_partials).  Weak scaling; value = N*2^log2n / max-over-ranks time.
  extra      "commit": batched blob commitments (config 3: blobs x 4096 coefficients, sharded by blob, no
             collective) reported as commits/s in the same JSON line.
Reference arm (--impl reference): the UNMODIFIED reference library (oracle/_ref/kzgref-bls12381, built from
/root/reference by oracle/build_ref.sh) runs trusted_setup::create_commit == polyeval_G1 on the host cores:
one process per host thread (the reference itself is single-threaded on this path), each a bounded sample
of the same workload (4096 of the points).  Rank 0 only.

The oracle (oracle/) is used here ONLY for the cpu_baseline / reference legs, never on the product path.
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
def reference_points_per_s(sample_points, procs, reps, workdir, srs_path, seed=1):
    """Time the reference's create_commit (polyeval_G1, src/trusted_setup.cpp:149-174) on `procs` host
    processes at once, each over `sample_points` terms.  Returns (points/s, ms per rep of the slowest)."""
    import numpy as np
    paths = []
    for p in range(procs):
        path = os.path.join(workdir, "coeffs%d.bin" % p)
        random_scalars(sample_points, seed * 1000 + p).view(np.uint8).tofile(path)
        paths.append(path)
    t0 = time.time()
    ps = [subprocess.Popen([REF_BIN, "time", srs_path, path, "commit", str(reps)], stdout=subprocess.PIPE, text=True)
          for path in paths]
    ms = [float(p.communicate()[0].strip()) for p in ps]
    wall = time.time() - t0
    if any(p.returncode for p in ps):
        raise RuntimeError("reference harness failed")
    worst = max(ms)
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
    n = 1 << args.log2n
    with tempfile.TemporaryDirectory() as td:
        srs = make_reference_srs(sample + 1, td)
        for w in range(args.warmup):
            reference_points_per_s(256, cores, 1, td, srs, seed=90 + w)
        t_total, pts = 0.0, 0
        for k in range(args.steps):
            v, worst_ms, _ = reference_points_per_s(sample, cores, 1, td, srs, seed=k + 1)
            t_total += worst_ms
            pts += cores * sample
    value = pts / (t_total * 1e-3)
    line = {
        "impl": "reference", "metric": "g1_msm_points_per_s", "value": value, "unit": "points/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32 limbs (reference: 58-bit limbs in int64)", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": "points/s", "cores": cores, "kind": "reference",
                         "sample": "%d host processes x create_commit over %d of the 2^%d terms per step (reference MSM is "
                                   "strictly per-term, so points/s is size-independent); harness-level parallelism, the "
                                   "reference itself is single-threaded here" % (cores, sample, args.log2n)},
        "e2e": {"value": value, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
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
    dist = None
    # Keep stdout to the one JSON line: libraries (NCCL prints its version banner to fd 1) write to stderr instead.
    sys.stdout.flush()
    json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    n = 1 << args.log2n
    ctx = kzg_b200.Context(CURVE, local)
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", local))
    ctx.set_option("setup_g2", 0)
    ctx.set_option("profile", 1)
    ctx.set_option("precompute", args.precompute)
    ctx.set_option("window_bits", args.window_bits)
    # Point-range split: rank g owns n points and its own n scalars of the N*n-term sum.  The synthetic bases of
    # every rank are the powers [s^i]G1, i < n (same cost as any other n points); the scalars differ per rank.
    ctx.setup_powers(SECRET, n, want_bytes=False)
    h_scalars = torch.from_numpy(random_scalars(n, 0x4B5A47 + rank).view(np.uint8).reshape(-1)).pin_memory()
    d_scalars = h_scalars.cuda(non_blocking=False)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    launches0 = ctx.launch_count

    def fold_partials():
        """N > 1: all-gather the projective partial sums (NCCL) and fold them on every rank."""
        return sharding.gather_partials_device(dist, ctx, torch.device("cuda", local))

    def step_resident():
        ctx.set_option("finish_affine", 0 if world > 1 else 1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        flush.zero_()
        torch.cuda.synchronize()
        e0.record(stream)
        ctx.msm_g1_launch(d_scalars, n, 1, 0)
        if world > 1:  # MSM -> NCCL all-gather -> fold, all ordered on the context's stream and inside the events
            with torch.cuda.stream(stream):
                fold_partials()  # (the per-rank MSM skipped its own affine normalisation: finish_affine = 0)
        e1.record(stream)
        e1.synchronize()
        return e0.elapsed_time(e1)

    d_stage = torch.empty_like(d_scalars) if world > 1 else None

    def step_e2e():
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if world == 1:   # the call polyeval_G1 makes in the drop-in: host scalars in, affine bytes out
            ctx.set_option("finish_affine", 1)
            xy, inf = ctx.msm_g1(h_scalars, n, 0)
        else:            # host scalars -> device, projective partial, NCCL all-gather, fold + normalise, bytes to host
            ctx.set_option("finish_affine", 0)
            with torch.cuda.stream(stream):
                d_stage.copy_(h_scalars)  # (blocking: a non_blocking copy would tie the pinned buffer's lifetime to this stream)
                ctx.msm_g1_launch(d_stage, n, 1, 0)
                xy, inf = fold_partials()
        return (time.perf_counter() - t0) * 1e3, xy

    for _ in range(max(args.warmup, 3)):
        step_resident()
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    if sampler:
        sampler.mark_start()
    l_before = ctx.launch_count
    step_ms, phase_log = [], []
    for _ in range(args.steps):
        step_ms.append(step_resident())
        phase_log.append(ctx.phase_times())  # CUDA events recorded on the context's stream around each pipeline phase
    barrier()
    gpu_launches = ctx.launch_count - l_before
    phases = {k: sum(p[k] for p in phase_log) / len(phase_log) for k in phase_log[0]}
    stats = ctx.msm_stats()
    e2e_ms, xy = [], None
    step_e2e()
    for _ in range(args.steps):
        ms, xy = step_e2e()
        e2e_ms.append(ms)
    barrier()
    if sampler:
        sampler.mark_stop()
    xy_resident, _ = (ctx.msm_fetch(1) if world == 1 else (xy, None))
    if world == 1:
        ctx.set_option("finish_affine", 1)
        ctx.msm_g1_launch(d_scalars, n, 1, 0)
        xy_resident, _ = ctx.msm_fetch(1)
        assert xy_resident == xy, "resident and end-to-end paths disagree"

    # batched blob commitments (config 3), sharded by blob across ranks, no collective
    commit = None
    if args.blobs > 0:
        nb, nc = args.blobs, 4096
        ctx.set_option("window_bits", 0)
        ctx.set_option("finish_affine", 1)
        ctx.setup_powers(SECRET, nc + 1, want_bytes=False)
        hb = torch.from_numpy(random_scalars(nb * nc, 77 + rank).view(np.uint8).reshape(-1)).pin_memory()
        db = hb.cuda()
        times = []
        for it in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            flush.zero_()
            torch.cuda.synchronize()
            e0.record(stream)
            ctx.msm_g1_launch(db, nc, nb, 0)
            e1.record(stream)
            e1.synchronize()
            times.append(e0.elapsed_time(e1))
        ctx.msm_g1_batch(hb, nc, nb)  # untimed: grows the staging buffer for this batch size
        t0 = time.perf_counter()
        ctx.msm_g1_batch(hb, nc, nb)
        t_e2e = (time.perf_counter() - t0) * 1e3
        # batched single-point proofs (create_proof(poly, off, 1) for every blob): quotient + MSM on device
        roots = [int(i % 4096) for i in range(nb)]
        ctx.prove_batch(hb, nc, roots, 1)
        t0 = time.perf_counter()
        ctx.prove_batch(hb, nc, roots, 1)
        t_prove = (time.perf_counter() - t0) * 1e3
        commit = {"blobs_per_gpu": nb, "coeffs": nc, "ms_resident": min(times[1:]), "ms_e2e": t_e2e, "ms_prove_e2e": t_prove}

    def allmax(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    total_ms = allmax(sum(step_ms))
    total_e2e_ms = allmax(sum(e2e_ms))
    if commit:
        commit["ms_resident"] = allmax(commit["ms_resident"])
        commit["ms_e2e"] = allmax(commit["ms_e2e"])
        commit["ms_prove_e2e"] = allmax(commit["ms_prove_e2e"])

    if rank == 0:
        points = n * world
        ms_per_step = total_ms / args.steps
        value = points / (ms_per_step * 1e-3)
        e2e_value = points / (total_e2e_ms / args.steps * 1e-3)
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
        line = {
            "metric": "g1_msm_points_per_s", "value": value, "unit": "points/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32 limbs (Montgomery, 12x32-bit Fp / 8x32-bit Fr)", "data": "synthetic",
            "config": workload_config(args),
            "e2e": {"value": e2e_value, "unit": "points/s", "h2d_bytes_per_step": n * 32, "d2h_bytes_per_step": 96 + 4,
                    "ms_per_step": total_e2e_ms / args.steps},
            "gpu_launches": gpu_launches,
            "roofline": {"bound": "int32-imad", "kernel": "msm_accumulate", "achieved": achieved / 1e12, "peak": peak / 1e12,
                         "unit": "TIMAD/s", "frac": achieved / peak,
                         # dram__bytes_read.sum + dram__bytes_write.sum of one msm_accumulate launch at this workload from
                         # the ncu --set full capture summarised in profiles/r1i_accumulate_ncu.md (2.734 + 0.304 GB)
                         "traffic": 3.038e9 if (args.log2n == 20 and args.precompute == 1 and not args.window_bits) else None,
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
        if commit:
            line["commit"] = {"metric": "blob_commits_per_s", "value": commit["blobs_per_gpu"] * world / (commit["ms_resident"] * 1e-3),
                              "e2e": commit["blobs_per_gpu"] * world / (commit["ms_e2e"] * 1e-3), "unit": "commits/s",
                              "blobs": commit["blobs_per_gpu"] * world, "coeffs_per_blob": commit["coeffs"],
                              "ms_resident": commit["ms_resident"], "ms_e2e": commit["ms_e2e"],
                              "proofs_per_s_e2e": commit["blobs_per_gpu"] * world / (commit["ms_prove_e2e"] * 1e-3),
                              "ms_prove_e2e": commit["ms_prove_e2e"]}
        # CPU baseline: the unmodified reference on the host cores, bounded sample (N = 1 only)
        if world == 1 and not args.no_cpu and os.path.exists(REF_BIN):
            try:
                with tempfile.TemporaryDirectory() as td:
                    sample = args.ref_sample
                    srs = make_reference_srs(sample + 1, td)
                    v, worst_ms, wall = reference_points_per_s(sample, 1, 1, td, srs)
                    line["cpu_baseline"] = {"value": v, "unit": "points/s", "cores": 1, "kind": "reference",
                                            "sample": "reference create_commit (polyeval_G1) over %d of the 2^%d terms, 1 thread "
                                                      "(the reference MSM is single-threaded and strictly per-term): %.0f ms"
                                                      % (sample, args.log2n, worst_ms)}
            except Exception as e:  # the baseline leg must never sink the GPU numbers
                line["cpu_baseline"] = {"value": None, "unit": "points/s", "cores": 1, "kind": "reference", "sample": "failed: %r" % (e,)}
        json_out.write(json.dumps(line) + "\n")
        json_out.flush()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
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
                    help="blobs per GPU for the batched-commit line (0 = skip); 4096 at N=1 is BASELINE config 3 as named")
    ap.add_argument("--ref-sample", type=int, default=16384, help="terms per reference process per step")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.ref_sample == 16384:
            args.ref_sample = 4096
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
