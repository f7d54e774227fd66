"""Development probe: the powers-of-tau batch (kzg_gpu_setup_powers) timed on its own.
Wall-clock around the synchronous C-ABI call (it ends with a stream synchronise), after one warm-up call that also
builds the per-context fixed-base tables.  precompute = 0 isolates A8 (power chain + fixed-base comb + one
inversion per output); the per-SRS MSM window table is timed separately as the difference with precompute on.
Work figure: per index 32 mixed additions in G1 (10 Fp modmul) and in G2 (28) = 1216 Fp modmul,
600 IMAD-class slots each on BLS12-381; nominal pipe peak 64 slots/clk/SM x 148 SMs x 1965 MHz."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import kzg_b200

PEAK = 64 * 148 * 1965e6
IMAD = {"bls12381": 600, "bn254": 272, "bn158": 2 * (2 * 36 + 6)}
out = []
t0 = time.time()
for curve in ("bls12381", "bn254"):
    t0 = time.time()
    ctx = kzg_b200.Context(curve, 0)
    t_ctx = time.time() - t0
    ctx.set_option("precompute", 0)
    t0 = time.time()
    ctx.setup_powers(987654321987654321, 256, want_bytes=False)
    t_first = time.time() - t0
    for n in (4096, 65536, 1 << 20):
        row = {"curve": curve, "n": n, "ctx_create_ms": t_ctx * 1e3, "first_call_ms": t_first * 1e3}
        for g2 in (0, 1):
            ctx.set_option("setup_g2", g2)
            for want in (False, True):
                if want and n > 65536:
                    continue
                ts = []
                for rep in range(3):
                    t0 = time.time()
                    ctx.setup_powers(987654321987654321 + rep, n, want_bytes=want)
                    ts.append(time.time() - t0)
                key = ("g1g2" if (g2 or want) else "g1") + ("_bytes" if want else "")
                row[key + "_ms"] = min(ts) * 1e3
        modmul = 320 + 896
        row["g1g2_indices_per_s"] = n / (row["g1g2_ms"] * 1e-3)
        row["g1g2_frac_of_imad_peak"] = row["g1g2_indices_per_s"] * modmul * IMAD[curve] / PEAK
        row["g1_frac_of_imad_peak"] = n / (row["g1_ms"] * 1e-3) * 320 * IMAD[curve] / PEAK
        ctx.set_option("setup_g2", 0)
        ctx.set_option("precompute", 1)
        ts = []
        for rep in range(2):
            t0 = time.time()
            ctx.setup_powers(987654321987654321 + rep, n, want_bytes=False)
            ts.append(time.time() - t0)
        row["g1_with_msm_table_ms"] = min(ts) * 1e3
        ctx.set_option("precompute", 0)
        out.append(row)
        print(json.dumps(row), flush=True)
    ctx.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "probe_setup.json"), "w"), indent=1)
