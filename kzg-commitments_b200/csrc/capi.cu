// C ABI (include/kzg_gpu.h) over the CUDA kernels: one templated context per curve parameter set
// (config/curve_{BN158,BN254,BLS12381}/kzg_config.h in the reference select the same three sets at
// compile time; here all three live in one library and are selected at kzg_gpu_ctx_create).
#include "../../include/kzg_gpu.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "curve_constants.cuh"
#include "diag.cuh"
#include "msm.cuh"
#include "poly.cuh"
#include "setup.cuh"

using namespace kzgb200;

namespace {

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

int ilog2(uint64_t v) {
  int r = 0;
  while (v >>= 1) r++;
  return r;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
struct kzg_gpu_ctx {
  int curve = -1;
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string last_error;
  uint64_t launches = 0;
  std::mutex mu;  // the reference API is re-entrant for readers; serialise use of the shared scratch
  // options
  int opt_window_bits = 0, opt_precompute = 0, opt_task_chunk = 0, opt_reduce_chunk = 0;

  virtual ~kzg_gpu_ctx() {}
  virtual int srs_load(uint64_t n, const uint8_t* g1, const uint8_t* g2) = 0;
  virtual int setup_powers(const uint8_t* s_le, uint64_t n, uint8_t* g1_out, uint8_t* g2_out) = 0;
  virtual int msm(int group, const uint8_t* scalars, bool scalars_on_device, uint64_t count, uint64_t batch,
                  uint64_t base_offset, uint8_t* out, int* is_inf, bool fetch) = 0;
  virtual int fetch(uint8_t* out, int* is_inf) = 0;
  virtual int fetch_partial(uint32_t* words, size_t* n_words) = 0;
  virtual int sum_partials(const uint32_t* words, uint64_t n, uint8_t* out, int* is_inf) = 0;
  virtual int poly_quotient(const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* q_out) = 0;
  virtual int poly_eval(const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* y_out) = 0;
  virtual int prove(const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* out, int* is_inf) = 0;
  virtual int microbench(int kind, double* ops) = 0;
  virtual int selftest(int* failures) = 0;

  int cuda_fail(cudaError_t e, const char* where) {
    char buf[256];
    snprintf(buf, sizeof buf, "%s: %s", where, cudaGetErrorString(e));
    last_error = buf;
    return e == cudaErrorMemoryAllocation ? KZG_GPU_ERR_MEMORY : KZG_GPU_ERR_CUDA;
  }
};

#define CK(call)                                                  \
  do {                                                            \
    cudaError_t e__ = (call);                                     \
    if (e__ != cudaSuccess) return cuda_fail(e__, #call);         \
  } while (0)
#define LAUNCHED()                                                \
  do {                                                            \
    launches++;                                                   \
    cudaError_t e__ = cudaGetLastError();                         \
    if (e__ != cudaSuccess) return cuda_fail(e__, "kernel launch"); \
  } while (0)

template <class C>
struct Ctx final : kzg_gpu_ctx {
  using Fp = Field<typename C::FpP>;
  using Fp2 = Field2<typename C::FpP>;
  using Fr = Field<typename C::FrP>;
  static constexpr int M = C::MODBYTES;
  static constexpr int SB = C::SCALAR_BYTES;
  static constexpr int MAXW = 40;

  // resident SRS
  uint64_t srs_n = 0;
  bool have_g2 = false;
  DevBuf g1, g2;             // Affine<Fp>[n], Affine<Fp2>[n]
  DevBuf g1_table;           // precomputed window table (row 0 unused: g1 itself is row 0's content)
  int table_c = 0, table_W = 0;
  // fixed-base tables of the generators
  DevBuf fb1, fb2;
  bool fb_ready = false;
  // scratch
  DevBuf stage, scal, digits, counts, starts, task_first, cursor, sorted, partial, chunk_out, group_out, out_bytes,
      out_inf, out_xyzz, poly0, poly1, qbytes, powers;
  // description of the most recent launch, for fetch()
  uint64_t last_batch = 0;
  int last_group = 1;

  ~Ctx() override {
    cudaSetDevice(device);
    for (DevBuf* b : {&g1, &g2, &g1_table, &fb1, &fb2, &stage, &scal, &digits, &counts, &starts, &task_first, &cursor,
                      &sorted, &partial, &chunk_out, &group_out, &out_bytes, &out_inf, &out_xyzz, &poly0, &poly1,
                      &qbytes, &powers})
      b->release();
    if (stream) cudaStreamDestroy(stream);
  }

  int init() {
    CK(cudaSetDevice(device));
    CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    return KZG_GPU_OK;
  }

  // ------------------------------------------------------------------ SRS
  int build_table() {
    g1_table.release();
    table_c = table_W = 0;
    if (!opt_precompute || srs_n == 0) return KZG_GPU_OK;
    int c = opt_window_bits ? opt_window_bits : std::min(16, std::max(8, ilog2(srs_n) - 2));
    int W = (C::SCALAR_BITS + 1 + c - 1) / c;
    if (W > MAXW) return KZG_GPU_ERR_ARG;
    CK(g1_table.ensure((size_t)W * srs_n * sizeof(Affine<Fp>)));
    msm_precompute_table<Fp, MAXW><<<(unsigned)((srs_n + 127) / 128), 128, 0, stream>>>(
        srs_n, srs_n, c, W, g1.as<Affine<Fp>>(), g1_table.as<Affine<Fp>>());
    LAUNCHED();
    CK(cudaStreamSynchronize(stream));
    table_c = c;
    table_W = W;
    return KZG_GPU_OK;
  }

  int srs_load(uint64_t n, const uint8_t* h_g1, const uint8_t* h_g2) override {
    if (!h_g1 || n == 0) return KZG_GPU_ERR_ARG;
    CK(cudaSetDevice(device));
    CK(g1.ensure(n * sizeof(Affine<Fp>)));
    CK(stage.ensure(n * (size_t)(4 * M)));
    CK(cudaMemcpyAsync(stage.p, h_g1, n * (size_t)(2 * M), cudaMemcpyHostToDevice, stream));
    points_from_bytes<Fp><<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(n, stage.as<uint8_t>(), M, g1.as<Affine<Fp>>());
    LAUNCHED();
    CK(cudaStreamSynchronize(stream));
    have_g2 = false;
    if (h_g2) {
      CK(g2.ensure(n * sizeof(Affine<Fp2>)));
      CK(cudaMemcpyAsync(stage.p, h_g2, n * (size_t)(4 * M), cudaMemcpyHostToDevice, stream));
      points_from_bytes<Fp2><<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(n, stage.as<uint8_t>(), M, g2.as<Affine<Fp2>>());
      LAUNCHED();
      CK(cudaStreamSynchronize(stream));
      have_g2 = true;
    }
    srs_n = n;
    return build_table();
  }
