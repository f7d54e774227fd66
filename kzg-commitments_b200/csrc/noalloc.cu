// kzg_gpu_poly_eval_range_noalloc — an entry point that allocates NOTHING (no cudaMalloc / cudaFree / allocating
// library call), for callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that
// forbid it.  It has no counterpart in the reference.
//
// It is the evaluation path of kzg_gpu_poly_eval_range (ctx.cuh poly_eval: fr_from_bytes -> poly_eval_range_kernel ->
// fr_to_bytes, i.e. evaluate_polynomial_points, src/util.cpp:186-211) with the two scratch buffers that path keeps in
// cudaMalloc'ed memory — the coefficients in Montgomery form and the Montgomery-form results — placed in ONE file-scope
// static __device__ array of fixed size instead.  Inputs and outputs are the caller's DEVICE buffers.
//
//   * Limit: n <= KZG_GPU_NOALLOC_MAX_COEFFS (65536) coefficients, k <= KZG_GPU_NOALLOC_MAX_POINTS (65536) points;
//     larger sizes are rejected with KZG_GPU_ERR_ARG and a message in kzg_gpu_last_error.  The array is 4 MiB and every
//     GPU that loads the library holds it until the process exits.
//   * In HOST code the array's name is only a host-side stand-in for the symbol, not a device address: the address is
//     looked up with cudaGetSymbolAddress once per device, with that device current, and that pointer is what the
//     kernels receive.
//   * Every call on a device shares the array, and it keeps the last call's contents.  Nothing in it has to start at
//     zero: a call overwrites coefficients [0, n) and results [0, k) before it reads them and reads nothing beyond.
//     Calls that OVERLAP on one GPU (two contexts / two streams) race on it — serialise them.
//   * The work is enqueued on the context's stream (kzg_gpu_ctx_stream) and the call returns without waiting; the
//     output is valid once that stream has reached the end of the three kernels.  The first call per device also loads
//     the kernels' module (CUDA loads lazily): make it once before starting a graph capture.
#include <mutex>

#include "ctx.cuh"

using namespace kzgb200;

namespace {

constexpr uint64_t MAX_COEFFS = KZG_GPU_NOALLOC_MAX_COEFFS;
constexpr uint64_t MAX_POINTS = KZG_GPU_NOALLOC_MAX_POINTS;
constexpr size_t FR_BYTES_MAX = 32;  // sizeof(Field<FrP>) is 32 (BN254, BLS12-381) or 24 (BN158)
constexpr size_t COEFF_REGION = MAX_COEFFS * FR_BYTES_MAX;  // results start here (a multiple of 256)

static __device__ __align__(256) unsigned char g_noalloc_scratch[(MAX_COEFFS + MAX_POINTS) * FR_BYTES_MAX];

constexpr int MAX_DEVICES = 64;
std::mutex g_setup_mu;
unsigned char* g_scratch_on_device[MAX_DEVICES] = {};

// device address of the scratch array on `device` (which must be current); looked up once per device
cudaError_t scratch_address(int device, unsigned char** out) {
  std::lock_guard<std::mutex> lock(g_setup_mu);
  if (!g_scratch_on_device[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_scratch_on_device[device] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_on_device[device];
  return cudaSuccess;
}

template <class C>
cudaError_t enqueue(unsigned char* scratch, const uint8_t* d_coeffs, uint64_t n, long long first_root, uint64_t k,
                    uint8_t* d_y_out, cudaStream_t stream) {
  using Fr = Field<typename C::FrP>;
  static_assert(sizeof(Fr) <= FR_BYTES_MAX, "scratch is sized for 32-byte field elements");
  Fr* coeffs = reinterpret_cast<Fr*>(scratch);
  Fr* ys = reinterpret_cast<Fr*>(scratch + COEFF_REGION);
  constexpr int ET = 128;
  fr_from_bytes<C><<<(unsigned)std::min<uint64_t>((n + 127) / 128, 148 * 8), 128, 0, stream>>>(n, d_coeffs, coeffs);
  poly_eval_range_kernel<C, ET><<<(unsigned)std::min<uint64_t>(k, 148 * 16), ET, 0, stream>>>(n, first_root, k, coeffs, ys);
  fr_to_bytes<C><<<(unsigned)std::min<uint64_t>((k + 127) / 128, 148 * 8), 128, 0, stream>>>(k, ys, d_y_out);
  return cudaGetLastError();
}

}  // namespace

extern "C" int kzg_gpu_poly_eval_range_noalloc(kzg_gpu_ctx* ctx, const void* d_coeffs, uint64_t n, int64_t first_root,
                                               uint64_t k, void* d_y_out) {
  if (!ctx) return KZG_GPU_ERR_ARG;
  std::lock_guard<std::recursive_mutex> lock(ctx->mu);
  ctx->last_error.clear();
  if (!d_coeffs || !d_y_out || n == 0 || k == 0) {
    ctx->last_error = "poly_eval_range_noalloc: null buffer or empty size";
    return KZG_GPU_ERR_ARG;
  }
  if (n > MAX_COEFFS || k > MAX_POINTS) {
    char buf[200];
    snprintf(buf, sizeof buf, "poly_eval_range_noalloc: n = %llu / k = %llu exceed the static scratch (at most %llu coefficients, %llu points)",
             (unsigned long long)n, (unsigned long long)k, (unsigned long long)MAX_COEFFS, (unsigned long long)MAX_POINTS);
    ctx->last_error = buf;
    return KZG_GPU_ERR_ARG;
  }
  if (ctx->device < 0 || ctx->device >= MAX_DEVICES) return KZG_GPU_ERR_ARG;
  int previous = -1;
  cudaError_t e = cudaGetDevice(&previous);
  if (e == cudaSuccess && previous != ctx->device) e = cudaSetDevice(ctx->device);
  unsigned char* scratch = nullptr;
  if (e == cudaSuccess) e = scratch_address(ctx->device, &scratch);
  if (e == cudaSuccess) {
    const uint8_t* in = static_cast<const uint8_t*>(d_coeffs);
    uint8_t* out = static_cast<uint8_t*>(d_y_out);
    switch (ctx->curve) {
      case KZG_GPU_BN158: e = enqueue<Curve_BN158>(scratch, in, n, (long long)first_root, k, out, ctx->stream); break;
      case KZG_GPU_BN254: e = enqueue<Curve_BN254>(scratch, in, n, (long long)first_root, k, out, ctx->stream); break;
      case KZG_GPU_BLS12381: e = enqueue<Curve_BLS12381>(scratch, in, n, (long long)first_root, k, out, ctx->stream); break;
      default: e = cudaErrorInvalidValue; break;
    }
    if (e == cudaSuccess) ctx->launches += 3;
  }
  if (previous >= 0 && previous != ctx->device) cudaSetDevice(previous);
  if (e != cudaSuccess) return ctx->cuda_fail(e, "poly_eval_range_noalloc");
  return KZG_GPU_OK;
}
