/* kzg_gpu.h — C ABI of the B200 (sm_100a) hot path for uncommitted6453/kzg-commitments.
 *
 * This is the ONLY interface the host side (the kzg.h drop-in in src/, the Python mirror in
 * kzg-commitments_b200/, tests and bench.py) links against.  Plain pointers and sizes, no C++ or
 * torch types.  Every function returns 0 on success and a negative kzg_gpu_status otherwise; nothing
 * throws or aborts.  There is NO CPU fallback behind any entry point: without a CUDA device
 * kzg_gpu_ctx_create fails with KZG_GPU_ERR_CUDA.
 *
 * Reference interfaces replaced (paths relative to the reference repository):
 *   kzg_gpu_msm_g1 / _batch        trusted_setup::polyeval_G1     src/trusted_setup.cpp:149-174
 *                                  (called by create_commit :137-142, create_proof :227, verify_proof :246)
 *   kzg_gpu_msm_g2                 trusted_setup::polyeval_G2     src/trusted_setup.cpp:176-201
 *   kzg_gpu_poly_quotient          ZZ_pX q = (P - I) / Z          src/trusted_setup.cpp:218-225
 *   kzg_gpu_poly_eval_range        evaluate_polynomial_points     src/util.cpp:186-211
 *   kzg_gpu_prove                  create_proof body              src/trusted_setup.cpp:214-228
 *   kzg_gpu_interpolate_range      polyfit / linear_roots_and_polyfit   src/util.cpp:172-184 (consecutive nodes)
 *   kzg_gpu_commit_bytes           blob::from_bytes + poly::from_blob + create_commit   src/blob.cpp:19-48
 *   kzg_gpu_setup_powers           trusted_setup(int) + generate_elements_range
 *                                                                 src/trusted_setup.cpp:21-74, 123-135
 *   kzg_gpu_srs_load               the _G1/_G2 vectors a trusted_setup holds (src/kzg.h:183-185),
 *                                  in the octet-body layout of export_setup (src/trusted_setup.cpp:256-287)
 *
 * Byte formats (identical to what the reference itself produces, so no translation layer is needed):
 *   scalar / coefficient  SB = CURVE_ORDER_BYTES little-endian bytes = BytesFromZZ(buf, rep(c), SB)
 *                         (src/util.cpp:31-43).  Values >= r are reduced mod r like PAIR_G1mul does.
 *   G1 point              x | y, each M = MODBYTES_CURVE big-endian canonical bytes = the body of
 *                         ECP_toOctet(.., compress=false) after the 0x04 tag (miracl-core/cpp/ecp.cpp:445-491).
 *   G2 point              x.b | x.a | y.b | y.a (imaginary part first; ecp2.cpp:209-215, fp2.cpp:83-87).
 *   infinity              all-zero coordinates (what ECP_toOctet leaves for the point at infinity),
 *                         additionally flagged through the is_inf outputs.
 *   M / SB:  BN158 20/20, BN254 32/32, BLS12381 48/32.
 */
#ifndef KZG_GPU_H
#define KZG_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct kzg_gpu_ctx kzg_gpu_ctx; /* opaque: owns device memory, stream, tables */

enum kzg_gpu_curve { KZG_GPU_BN158 = 0, KZG_GPU_BN254 = 1, KZG_GPU_BLS12381 = 2 };

enum kzg_gpu_status {
  KZG_GPU_OK = 0,
  KZG_GPU_ERR_ARG = -1,    /* bad argument (null pointer, range outside the loaded SRS, unknown curve) */
  KZG_GPU_ERR_CUDA = -2,   /* CUDA runtime error (no device, launch failure); see kzg_gpu_last_error */
  KZG_GPU_ERR_MEMORY = -3, /* device or pinned allocation failed */
  KZG_GPU_ERR_STATE = -4   /* call order (e.g. MSM before an SRS is resident) */
};

const char* kzg_gpu_strerror(int status);
/* error text of the most recent call on this context (empty string after a successful call); the pointer stays valid
 * until the next call on the same context */
const char* kzg_gpu_last_error(const kzg_gpu_ctx* ctx);

/* sizes for a curve: M (MODBYTES_CURVE) and SB (CURVE_ORDER_BYTES); < 0 for an unknown curve */
int kzg_gpu_modbytes(int curve);
int kzg_gpu_scalar_bytes(int curve);

/* One context per (curve, device).  `device` is a CUDA ordinal. */
int kzg_gpu_ctx_create(kzg_gpu_ctx** out, int curve, int device);
void kzg_gpu_ctx_destroy(kzg_gpu_ctx* ctx);

/* Tunables (mostly for bench/tests): "window_bits" (0 = auto), "precompute" (0: none, 1: build the per-SRS
 * G1 window table T[w][i] = 2^(c*w) G_i at srs_load / setup_powers, 2: also for G2), "profile" (record
 * per-phase CUDA events, see kzg_gpu_phase_times), "task_chunk", "reduce_chunk", "setup_g2" (0: setup_powers with a NULL
 * g2_out generates a G1-only SRS), "team" (1: latency-bound reduction levels on 4-lane teams), "long_division"
 * (many-root quotient of one polynomial: 0 systolic, 1 long division on one CTA, 2 on an 8-CTA cluster,
 * 3 = default: power-series inversion with truncated products over the whole GPU), "affine_levels" (batched-affine
 * pre-reduction levels, -1 = auto = off), "finish_affine" (0: an MSM leaves only its projective result for
 * kzg_gpu_msm_partial_device — the point-range split folds and normalises once, after the exchange).
 * "tail" (1: the sort and the narrow end of the reduction run in two persistent cooperative kernels), "tail_proj" (1: that
 * end uses complete projective formulas on 8-lane teams for G1; 2: also for G2, on whole-warp teams), "table_pad" (experiment: 128-byte window-table entries),
 * "direct_table_gb" (0 = off: GiB of device memory a DIRECT-LOOKUP table of the next SRS may take — every multiple
 * d 2^(cw) G_i a digit can select, csrc/msm_direct.cuh; 4097 points: 32 GiB at c = 13, 61 at 14, 110 at 15 — G1 MSMs then
 * skip sort, buckets and weighted reduction: for services that commit 4096-element blobs against one small SRS),
 * "table_ladder" (1: a large SRS also gets window tables of its first 8192 and 131072 points with window widths that suit
 * those sizes, used by MSMs that fit — a small commitment does not walk the bucket reduction of the big table),
 * "barrier_sleep" (ns between polls of the grid barrier).
 * Takes effect at the next srs_load / setup_powers / MSM. */
int kzg_gpu_set_option(kzg_gpu_ctx* ctx, const char* name, long long value);

/* Make an SRS resident.  g1: n * 2M bytes.  g2: n * 4M bytes or NULL (G1-only use). */
int kzg_gpu_srs_load(kzg_gpu_ctx* ctx, uint64_t n, const uint8_t* g1, const uint8_t* g2);

/* [s^i]G1 (n * 2M bytes) and [s^i]G2 (n * 4M bytes) for i < n from the secret s (SB little-endian
 * bytes); either output may be NULL.  The generated SRS is left resident (as after srs_load). */
int kzg_gpu_setup_powers(kzg_gpu_ctx* ctx, const uint8_t* s_le, uint64_t n, uint8_t* g1_out, uint8_t* g2_out);
/* The slice [first_index, first_index + n) of the same SRS: [s^(first_index + i)]G1 / G2, i < n, left resident as
 * points 0..n-1 of this context.  This is generate_elements_range(start, end, ..) of the reference
 * (src/trusted_setup.cpp:123-135: its std::thread fan-out over index ranges) with the device in place of a thread,
 * and what the point-range split of one large MSM across GPUs loads on device g.
 * The secret and its powers are zeroed in device memory before either call returns. */
int kzg_gpu_setup_powers_range(kzg_gpu_ctx* ctx, const uint8_t* s_le, uint64_t first_index, uint64_t n, uint8_t* g1_out,
                               uint8_t* g2_out);

/* out = sum_{i<count} scalars[i] * G1[base_offset + i]   (host buffers) */
int kzg_gpu_msm_g1(kzg_gpu_ctx* ctx, const uint8_t* scalars, uint64_t count, uint64_t base_offset,
                   uint8_t* out_xy /* 2M */, int* is_inf);
int kzg_gpu_msm_g2(kzg_gpu_ctx* ctx, const uint8_t* scalars, uint64_t count, uint64_t base_offset,
                   uint8_t* out_xy /* 4M */, int* is_inf);
/* `batch` independent MSMs over the same bases G1[0..count): scalars is batch*count*SB bytes,
 * out is batch*2M bytes, is_inf is batch ints. */
int kzg_gpu_msm_g1_batch(kzg_gpu_ctx* ctx, const uint8_t* scalars, uint64_t count, uint64_t batch,
                         uint8_t* out_xy, int* is_inf);

/* Device-resident variants: `d_scalars` is a DEVICE pointer (same byte layout).  _launch only
 * enqueues on the context stream (result stays on device); _fetch waits and copies the results of
 * the most recent launch.  Used for kernel-only timing and for pipelines that produce scalars on device. */
int kzg_gpu_msm_g1_launch(kzg_gpu_ctx* ctx, const void* d_scalars, uint64_t count, uint64_t batch, uint64_t base_offset);
int kzg_gpu_msm_fetch(kzg_gpu_ctx* ctx, uint8_t* out_xy, int* is_inf);
/* projective (XYZZ, Montgomery limbs) result of the most recent single launch, for the multi-GPU
 * point-range split: each rank fetches 4*L words, ranks exchange them (NCCL all-gather / host) and
 * one rank folds them with kzg_gpu_g1_sum_partials. */
int kzg_gpu_msm_fetch_partial(kzg_gpu_ctx* ctx, uint32_t* xyzz_words, size_t* n_words);
int kzg_gpu_g1_sum_partials(kzg_gpu_ctx* ctx, const uint32_t* xyzz_words, uint64_t n_partials,
                            uint8_t* out_xy /* 2M */, int* is_inf);

/* Device-resident forms of the two calls above for an NCCL exchange without host round trips: the device
 * address and byte size of the most recent single launch's projective result, and the fold of n_partials
 * such records lying contiguously in DEVICE memory (e.g. the output of ncclAllGather). */
int kzg_gpu_msm_partial_device(kzg_gpu_ctx* ctx, void** d_ptr, size_t* bytes);
int kzg_gpu_g1_sum_partials_device(kzg_gpu_ctx* ctx, const void* d_xyzz, uint64_t n_partials,
                                   uint8_t* out_xy /* 2M */, int* is_inf);

/* q = P div prod_{j<k} (x - (first_root + j)): coeffs n*SB bytes in, (n-k)*SB bytes out (k < n). */
int kzg_gpu_poly_quotient(kzg_gpu_ctx* ctx, const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k,
                          uint8_t* q_out);
/* y_j = P(first_root + j), j < k: k*SB bytes out. */
int kzg_gpu_poly_eval_range(kzg_gpu_ctx* ctx, const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k,
                            uint8_t* y_out);
/* create_proof: quotient and G1 MSM chained on device (the quotient never visits the host).
 * k >= n gives the zero quotient, i.e. the point at infinity (benchmark "Proofs: 4096" row). */
int kzg_gpu_prove(kzg_gpu_ctx* ctx, const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k,
                  uint8_t* out_xy /* 2M */, int* is_inf);

/* `batch` independent create_proof calls in one: polynomial b = coeffs + b*n*SB (n coefficients each, zero-padded),
 * opened at the k consecutive points first_roots[b] .. first_roots[b]+k-1.  out_xy: batch*2M bytes. */
int kzg_gpu_prove_batch(kzg_gpu_ctx* ctx, const uint8_t* coeffs, uint64_t n, const int64_t* first_roots, uint64_t k,
                        uint64_t batch, uint8_t* out_xy, int* is_inf);

/* Interpolation through the k consecutive nodes x_i = first_x + i (what every kzg::blob holds, src/blob.cpp:20-48):
 * replaces polyfit (src/util.cpp:180-184, used by poly::from_blob src/poly.cpp:4-6) and linear_roots_and_polyfit
 * (src/util.cpp:172-178, used by verify_proof src/trusted_setup.cpp:239).  ys: k*SB bytes; coeffs_out: the k
 * coefficients of the interpolant I (zero-padded, low degree first) or NULL; vanishing_out: the k+1 coefficients of
 * Z = prod (x - x_i) or NULL. */
int kzg_gpu_interpolate_range(kzg_gpu_ctx* ctx, const uint8_t* ys, int64_t first_x, uint64_t k, uint8_t* coeffs_out,
                              uint8_t* vanishing_out);

/* Document bytes to commitment without leaving the device: blob::from_bytes (src/blob.cpp:19-48: chunk i of chunk_size
 * bytes, zero-padded, little-endian, at node chunk_offset + i) -> poly::from_blob (src/poly.cpp:4-6) -> create_commit
 * (src/trusted_setup.cpp:137-147).  chunk_size in 1..SB-1 (MAX_CHUNK_BYTES, src/kzg.h:31) and byte_length a multiple of
 * it, else KZG_GPU_ERR_ARG (the reference throws invalid_argument); byte_length / chunk_size < SRS size (the bound of
 * create_commit, src/trusted_setup.cpp:138).
 * coeffs_out: NULL or byte_length/chunk_size * SB bytes receiving the polynomial (for later create_proof calls). */
int kzg_gpu_commit_bytes(kzg_gpu_ctx* ctx, const uint8_t* bytes, uint64_t byte_length, int chunk_size, int64_t chunk_offset,
                         uint8_t* coeffs_out, uint8_t* out_xy /* 2M */, int* is_inf);

/* The context's CUDA stream (cudaStream_t) so a caller can time with events on it. */
void* kzg_gpu_ctx_stream(kzg_gpu_ctx* ctx);
/* kernels launched by this context since creation (bench.py's gpu_launches) */
uint64_t kzg_gpu_launch_count(const kzg_gpu_ctx* ctx);

/* With the "profile" option set: CUDA-event durations (ms) of the most recent MSM pass, ms[0..3] =
 * digits+sort+tasks, bucket accumulation, bucket reduction, finish (affine + bytes).  n >= 4. */
int kzg_gpu_phase_times(kzg_gpu_ctx* ctx, float* ms, int n);

/* Shape of the most recent MSM pass, out[0..5] = records (non-trivial window digits incl. zeros), buckets, tasks,
 * most tasks in one bucket, window bits c, task chunk.  n >= 6. */
int kzg_gpu_msm_stats(kzg_gpu_ctx* ctx, uint64_t* out, int n);

/* Integer-pipe microbenchmarks used for the roofline denominator:
 * kind 0: independent mad.lo.cc/madc.hi.cc chains (-> IMAD.WIDE.U32.X), returns 32x32 products/s
 * kind 1: independent plain 32-bit IMAD (mad.lo.u32), returns instructions/s
 * kind 2: Fp Montgomery products/s of the context's curve (register-resident operands)
 * kind 3: Fr Montgomery products/s
 * kinds 4..7: LATENCY probes in the configuration of the reduction tail (one CTA of 128 threads per SM), returned as
 * steps per second of ONE dependent chain: 4 = grid barriers, 5 = 4-lane team additions (XYZZ, G1 field),
 * 6 = single-thread additions, 7 = dependent Fp Montgomery products, 8 = Fp inversions (batched divsteps),
 * 9 = Fp inversions by the binary extended GCD of round 1 (kept as a cross-check), 10 / 11 = G2 additions on 16-lane /
 * 4-lane XYZZ teams, 12 = G2 additions on whole-warp projective teams, 13 = G1 additions on 8-lane projective teams */
int kzg_gpu_microbench(kzg_gpu_ctx* ctx, int kind, double* ops_per_second);

/* Device self-test: field, Fp2 and group-law identities on random inputs, compared on device against
 * independent formulations (e.g. (a+b)^2 = a^2+2ab+b^2, P+Q-Q = P, [r]G = inf).  Returns the number of
 * failed checks in *failures. */
int kzg_gpu_selftest(kzg_gpu_ctx* ctx, int* failures);

/* kzg_gpu_poly_eval_range for callers that CANNOT ALLOCATE (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.
 * Same values as kzg_gpu_poly_eval_range (y_j = P(first_root + j), j < k), but d_coeffs (n*SB bytes) and d_y_out (k*SB
 * bytes) are DEVICE buffers of the caller, the call makes no cudaMalloc / cudaFree / allocating library call, and its
 * scratch is one fixed 4 MiB static __device__ array (held by every GPU that loads the library, until exit).
 * n > KZG_GPU_NOALLOC_MAX_COEFFS or k > KZG_GPU_NOALLOC_MAX_POINTS -> KZG_GPU_ERR_ARG with a message in kzg_gpu_last_error.
 * The work is enqueued on the context's stream (kzg_gpu_ctx_stream) and the call returns without waiting.  All calls on
 * one GPU share the scratch array (it keeps the last call's contents; nothing in it needs to start at zero): calls that
 * overlap on the GPU race — serialise them.  Make the first call per device before starting a graph capture (it looks
 * the array's address up and loads the kernels). */
#define KZG_GPU_NOALLOC_MAX_COEFFS 65536
#define KZG_GPU_NOALLOC_MAX_POINTS 65536
int kzg_gpu_poly_eval_range_noalloc(kzg_gpu_ctx* ctx, const void* d_coeffs, uint64_t n, int64_t first_root, uint64_t k,
                                    void* d_y_out);

/* ------------------------------------------------------------------------------------------------------------
 * Multi-GPU: a GROUP of devices of one box behind the same C ABI (SURVEY.md 8(b) "devices, n_devices" and 8(e)).
 *
 * Replaces the reference's only parallel construct on this path, the std::thread fan-out over index ranges
 * (src/trusted_setup.cpp:39-73), with one device context + one host worker thread per GPU:
 *   - independent units, no collective: the blobs of kzg_gpu_group_msm_g1_batch / _prove_batch are cut into
 *     contiguous ranges, one per device, each against that device's replica of the SRS; every device reads its
 *     slice straight from the caller's buffer and writes its results straight into the caller's arrays.
 *   - one exchange step: kzg_gpu_group_msm_g1 over at least `split_min` points is split by contiguous POINT
 *     RANGE; device g holds only its slice [g*n/G, (g+1)*n/G) of the SRS (with a window table sized for the slice),
 *     leaves one projective partial sum, and the <= 8 partials (4*L words each) are folded and normalised on device 0
 *     (kzg_gpu_g1_sum_partials).  One process drives all devices, so the exchange is G small D2H copies, not NCCL;
 *     the one-process-per-GPU form of the same split (NCCL all-gather of the partials) is in bench.py / sharding.py.
 * Calls on one group are serialised; different groups are independent.  No CPU fallback. */
typedef struct kzg_gpu_group kzg_gpu_group;

/* number of visible CUDA devices (0 when there is no driver or no device) */
int kzg_gpu_device_count(void);
/* devices == NULL: ordinals 0 .. n_devices-1 (n_devices <= 0: every visible device).  An ordinal may be listed more
 * than once (members sharing a device). */
int kzg_gpu_group_create(kzg_gpu_group** out, int curve, const int* devices, int n_devices);
void kzg_gpu_group_destroy(kzg_gpu_group* group);
int kzg_gpu_group_size(const kzg_gpu_group* group);
/* the context of member i holding the full (replicated) SRS: for the single-device entry points above */
kzg_gpu_ctx* kzg_gpu_group_ctx(kzg_gpu_group* group, int i);
const char* kzg_gpu_group_last_error(const kzg_gpu_group* group);
/* kzg_gpu_set_option on every member context; additionally "split_min" (points from which a single MSM is split
 * by point range, default 65536; must be set before the SRS is made resident). */
int kzg_gpu_group_set_option(kzg_gpu_group* group, const char* name, long long value);
/* as kzg_gpu_srs_load / kzg_gpu_setup_powers: the SRS becomes resident on every member (replica + range slice) */
int kzg_gpu_group_srs_load(kzg_gpu_group* group, uint64_t n, const uint8_t* g1, const uint8_t* g2);
int kzg_gpu_group_setup_powers(kzg_gpu_group* group, const uint8_t* s_le, uint64_t n, uint8_t* g1_out, uint8_t* g2_out);
/* as kzg_gpu_msm_g1 / _batch / kzg_gpu_prove_batch, sharded over the members as described above */
int kzg_gpu_group_msm_g1(kzg_gpu_group* group, const uint8_t* scalars, uint64_t count, uint64_t base_offset,
                         uint8_t* out_xy /* 2M */, int* is_inf);
int kzg_gpu_group_msm_g1_batch(kzg_gpu_group* group, const uint8_t* scalars, uint64_t count, uint64_t batch,
                               uint8_t* out_xy, int* is_inf);
int kzg_gpu_group_prove_batch(kzg_gpu_group* group, const uint8_t* coeffs, uint64_t n, const int64_t* first_roots,
                              uint64_t k, uint64_t batch, uint8_t* out_xy, int* is_inf);

#ifdef __cplusplus
}
#endif
#endif /* KZG_GPU_H */
