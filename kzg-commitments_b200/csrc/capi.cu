// extern "C" entry points of include/kzg_gpu.h; the per-curve contexts live in curve_*.cu.
#include "ctx.cuh"

using namespace kzgb200;

// ---------------------------------------------------------------------------------------------
extern "C" {

const char* kzg_gpu_strerror(int status) {
  switch (status) {
    case KZG_GPU_OK: return "ok";
    case KZG_GPU_ERR_ARG: return "invalid argument";
    case KZG_GPU_ERR_CUDA: return "CUDA error";
    case KZG_GPU_ERR_MEMORY: return "out of device memory";
    case KZG_GPU_ERR_STATE: return "invalid call order (no SRS resident?)";
    default: return "unknown status";
  }
}

const char* kzg_gpu_last_error(const kzg_gpu_ctx* ctx) { return ctx ? ctx->last_error.c_str() : ""; }

int kzg_gpu_modbytes(int curve) {
  switch (curve) {
    case KZG_GPU_BN158: return Curve_BN158::MODBYTES;
    case KZG_GPU_BN254: return Curve_BN254::MODBYTES;
    case KZG_GPU_BLS12381: return Curve_BLS12381::MODBYTES;
    default: return KZG_GPU_ERR_ARG;
  }
}
int kzg_gpu_scalar_bytes(int curve) {
  switch (curve) {
    case KZG_GPU_BN158: return Curve_BN158::SCALAR_BYTES;
    case KZG_GPU_BN254: return Curve_BN254::SCALAR_BYTES;
    case KZG_GPU_BLS12381: return Curve_BLS12381::SCALAR_BYTES;
    default: return KZG_GPU_ERR_ARG;
  }
}

int kzg_gpu_device_count(void) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess) { cudaGetLastError(); return 0; }
  return ndev;
}

int kzg_gpu_ctx_create(kzg_gpu_ctx** out, int curve, int device) {
  if (!out) return KZG_GPU_ERR_ARG;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return KZG_GPU_ERR_CUDA;  // no CPU fallback
  if (device < 0 || device >= ndev) return KZG_GPU_ERR_ARG;
  kzg_gpu_ctx* c = nullptr;
  int rc = KZG_GPU_OK;
  switch (curve) {
    case KZG_GPU_BN158: c = kzg_gpu_make_ctx_bn158(device, &rc); break;
    case KZG_GPU_BN254: c = kzg_gpu_make_ctx_bn254(device, &rc); break;
    case KZG_GPU_BLS12381: c = kzg_gpu_make_ctx_bls12381(device, &rc); break;
    default: return KZG_GPU_ERR_ARG;
  }
  if (!c) return rc ? rc : KZG_GPU_ERR_MEMORY;
  c->curve = curve;
  *out = c;
  return KZG_GPU_OK;
}

void kzg_gpu_ctx_destroy(kzg_gpu_ctx* ctx) { delete ctx; }

// Every entry point: serialise on the context, forget the previous call's error text, and leave a message behind for
// argument / call-order failures too (CUDA failures record theirs where they happen).
#define KZG_ENTER(ctx)                                        \
  if (!(ctx)) return KZG_GPU_ERR_ARG;                         \
  std::lock_guard<std::recursive_mutex> lock__((ctx)->mu);    \
  (ctx)->last_error.clear()
static int leave(kzg_gpu_ctx* ctx, int rc, const char* what) {
  if (rc != KZG_GPU_OK && ctx->last_error.empty())
    ctx->last_error = std::string(what) + ": " + kzg_gpu_strerror(rc);
  return rc;
}

int kzg_gpu_set_option(kzg_gpu_ctx* ctx, const char* name, long long value) {
  KZG_ENTER(ctx);
  if (!name) return leave(ctx, KZG_GPU_ERR_ARG, "set_option");
  std::string n(name);
  bool ok = true;
  if (n == "window_bits") { ok = value == 0 || (value >= 2 && value <= 22); if (ok) ctx->opt_window_bits = (int)value; }
  else if (n == "precompute") { ok = value >= 0 && value <= 2; if (ok) ctx->opt_precompute = (int)value; }
  else if (n == "profile") ctx->opt_profile = value ? 1 : 0;
  else if (n == "long_division") { ok = value >= 0 && value <= 3; if (ok) ctx->opt_long_division = (int)value; }
  else if (n == "finish_affine") ctx->opt_finish_affine = value ? 1 : 0;
  else if (n == "tail") { ok = value >= 0 && value <= 1; if (ok) ctx->opt_tail = (int)value; }
  else if (n == "barrier_sleep") { ok = value >= 0 && value <= 100000; if (ok) ctx->opt_barrier_sleep = (int)value; }
  else if (n == "tail_proj") { ok = value >= 0 && value <= 2; if (ok) ctx->opt_tail_proj = (int)value; }
  else if (n == "direct_table_gb") { ok = value >= 0 && value <= 1024; if (ok) ctx->opt_direct_table_gb = (int)value; }
  else if (n == "table_ladder") { ok = value >= 0 && value <= 1; if (ok) ctx->opt_table_ladder = (int)value; }
  else if (n == "table_pad") { ok = value >= 0 && value <= 1; if (ok) ctx->opt_table_pad = (int)value; }
  else if (n == "team") { ok = value >= 0 && value <= 1; if (ok) ctx->opt_team = (int)value; }
  else if (n == "affine_levels") { ok = value >= -1 && value <= 4; if (ok) ctx->opt_affine_levels = (int)value; }
  else if (n == "setup_g2") ctx->opt_setup_g2 = value ? 1 : 0;
  else if (n == "task_chunk") { ok = value >= 0 && value <= (1 << 20); if (ok) ctx->opt_task_chunk = (int)value; }
  else if (n == "reduce_chunk") { ok = value >= 0 && !(value & (value - 1)); if (ok) ctx->opt_reduce_chunk = (int)value; }
  else { ctx->last_error = "set_option: unknown option '" + n + "'"; return KZG_GPU_ERR_ARG; }
  if (!ok) { ctx->last_error = "set_option: value out of range for '" + n + "'"; return KZG_GPU_ERR_ARG; }
  return KZG_GPU_OK;
}

int kzg_gpu_srs_load(kzg_gpu_ctx* ctx, uint64_t n, const uint8_t* g1, const uint8_t* g2) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->srs_load(n, g1, g2), "srs_load");
}
int kzg_gpu_setup_powers(kzg_gpu_ctx* ctx, const uint8_t* s_le, uint64_t n, uint8_t* g1_out, uint8_t* g2_out) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->setup_powers(s_le, 0, n, g1_out, g2_out), "setup_powers");
}
int kzg_gpu_setup_powers_range(kzg_gpu_ctx* ctx, const uint8_t* s_le, uint64_t first_index, uint64_t n, uint8_t* g1_out,
                               uint8_t* g2_out) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->setup_powers(s_le, first_index, n, g1_out, g2_out), "setup_powers_range");
}
int kzg_gpu_msm_g1(kzg_gpu_ctx* ctx, const uint8_t* scalars, uint64_t count, uint64_t base_offset, uint8_t* out_xy, int* is_inf) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->msm(1, scalars, false, count, 1, base_offset, out_xy, is_inf, true), "msm_g1");
}
int kzg_gpu_msm_g2(kzg_gpu_ctx* ctx, const uint8_t* scalars, uint64_t count, uint64_t base_offset, uint8_t* out_xy, int* is_inf) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->msm(2, scalars, false, count, 1, base_offset, out_xy, is_inf, true), "msm_g2");
}
int kzg_gpu_msm_g1_batch(kzg_gpu_ctx* ctx, const uint8_t* scalars, uint64_t count, uint64_t batch, uint8_t* out_xy, int* is_inf) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->msm(1, scalars, false, count, batch, 0, out_xy, is_inf, true), "msm_g1_batch");
}
int kzg_gpu_msm_g1_launch(kzg_gpu_ctx* ctx, const void* d_scalars, uint64_t count, uint64_t batch, uint64_t base_offset) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->msm(1, static_cast<const uint8_t*>(d_scalars), true, count, batch, base_offset, nullptr, nullptr, false),
               "msm_g1_launch");
}
int kzg_gpu_msm_fetch(kzg_gpu_ctx* ctx, uint8_t* out_xy, int* is_inf) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->fetch(out_xy, is_inf), "msm_fetch");
}
int kzg_gpu_msm_fetch_partial(kzg_gpu_ctx* ctx, uint32_t* words, size_t* n_words) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->fetch_partial(words, n_words), "msm_fetch_partial");
}
int kzg_gpu_g1_sum_partials(kzg_gpu_ctx* ctx, const uint32_t* words, uint64_t n, uint8_t* out_xy, int* is_inf) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->sum_partials(words, n, out_xy, is_inf), "g1_sum_partials");
}
int kzg_gpu_msm_partial_device(kzg_gpu_ctx* ctx, void** d_ptr, size_t* bytes) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->partial_device(d_ptr, bytes), "msm_partial_device");
}
int kzg_gpu_g1_sum_partials_device(kzg_gpu_ctx* ctx, const void* d_words, uint64_t n, uint8_t* out_xy, int* is_inf) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->sum_partials_device(d_words, n, out_xy, is_inf), "g1_sum_partials_device");
}
int kzg_gpu_poly_quotient(kzg_gpu_ctx* ctx, const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* q_out) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->poly_quotient(coeffs, n, first_root, k, q_out), "poly_quotient");
}
int kzg_gpu_poly_eval_range(kzg_gpu_ctx* ctx, const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* y_out) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->poly_eval(coeffs, n, first_root, k, y_out), "poly_eval_range");
}
int kzg_gpu_prove(kzg_gpu_ctx* ctx, const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* out_xy, int* is_inf) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->prove(coeffs, n, first_root, k, out_xy, is_inf), "prove");
}
int kzg_gpu_prove_batch(kzg_gpu_ctx* ctx, const uint8_t* coeffs, uint64_t n, const int64_t* first_roots, uint64_t k,
                        uint64_t batch, uint8_t* out_xy, int* is_inf) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->prove_batch(coeffs, n, first_roots, k, batch, out_xy, is_inf), "prove_batch");
}
int kzg_gpu_interpolate_range(kzg_gpu_ctx* ctx, const uint8_t* ys, int64_t first_x, uint64_t k, uint8_t* coeffs_out,
                               uint8_t* vanishing_out) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->interpolate(ys, first_x, k, coeffs_out, vanishing_out), "interpolate_range");
}
int kzg_gpu_commit_bytes(kzg_gpu_ctx* ctx, const uint8_t* bytes, uint64_t byte_length, int chunk_size, int64_t chunk_offset,
                         uint8_t* coeffs_out, uint8_t* out_xy, int* is_inf) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->commit_bytes(bytes, byte_length, chunk_size, chunk_offset, coeffs_out, out_xy, is_inf), "commit_bytes");
}
void* kzg_gpu_ctx_stream(kzg_gpu_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
uint64_t kzg_gpu_launch_count(const kzg_gpu_ctx* ctx) { return ctx ? ctx->launches : 0; }
int kzg_gpu_microbench(kzg_gpu_ctx* ctx, int kind, double* ops) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->microbench(kind, ops), "microbench");
}
int kzg_gpu_phase_times(kzg_gpu_ctx* ctx, float* ms, int n) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->phase_times(ms, n), "phase_times");
}
int kzg_gpu_msm_stats(kzg_gpu_ctx* ctx, uint64_t* out, int n) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->msm_stats(out, n), "msm_stats");
}
int kzg_gpu_selftest(kzg_gpu_ctx* ctx, int* failures) {
  KZG_ENTER(ctx);
  return leave(ctx, ctx->selftest(failures), "selftest");
}

}  // extern "C"
