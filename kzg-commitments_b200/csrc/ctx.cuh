// Host-side context behind the C ABI (include/kzg_gpu.h): one templated context per curve parameter
// set.  config/curve_{BN158,BN254,BLS12381}/kzg_config.h in the reference select the same three sets
// at compile time; here each set is one translation unit (curve_*.cu, compiled in parallel) and all
// three live in one library, selected at kzg_gpu_ctx_create.
#pragma once
#include "../../include/kzg_gpu.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <string>
#include <type_traits>
#include <vector>

#include "curve_constants.cuh"
#include "diag.cuh"
#include "msm.cuh"
#include "msm_affine.cuh"
#include "msm_direct.cuh"
#include "msm_team.cuh"
#include "msm_tail.cuh"
#include "msm_tail_proj.cuh"
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
  cudaError_t ensure_exact(size_t bytes) {  // no head-room: tables of tens of GB
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e == cudaSuccess) cap = bytes;
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
  cudaStream_t stream = nullptr, copy_stream = nullptr;
  cudaEvent_t copied_ev[2] = {nullptr, nullptr}, consumed_ev[2] = {nullptr, nullptr};
  cudaEvent_t chunk_ev[4] = {nullptr, nullptr, nullptr, nullptr};
  int chunk_slices = 0;  // set by msm() right before enqueue_msm when the scalars are still arriving in chunks
  std::string last_error;
  uint64_t launches = 0;
  std::recursive_mutex mu;  // the reference API is re-entrant for readers; serialise use of the shared scratch
  // options
  int opt_window_bits = 0, opt_precompute = 2, opt_task_chunk = 0, opt_reduce_chunk = 0, opt_setup_g2 = 1,
      opt_profile = 0, opt_affine_levels = -1, opt_team = 1, opt_long_division = 3, opt_finish_affine = 1, opt_tail = 1, opt_table_pad = 0, opt_barrier_sleep = 64, opt_tail_proj = 1, opt_table_ladder = 1, opt_direct_table_gb = 0;
  cudaEvent_t phase_ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};

  virtual ~kzg_gpu_ctx() {}
  virtual int srs_load(uint64_t n, const uint8_t* g1, const uint8_t* g2) = 0;
  virtual int setup_powers(const uint8_t* s_le, uint64_t first_index, uint64_t n, uint8_t* g1_out, uint8_t* g2_out) = 0;
  virtual int msm(int group, const uint8_t* scalars, bool scalars_on_device, uint64_t count, uint64_t batch,
                  uint64_t base_offset, uint8_t* out, int* is_inf, bool fetch) = 0;
  virtual int fetch(uint8_t* out, int* is_inf) = 0;
  virtual int fetch_partial(uint32_t* words, size_t* n_words) = 0;
  virtual int sum_partials(const uint32_t* words, uint64_t n, uint8_t* out, int* is_inf) = 0;
  virtual int partial_device(void** d_ptr, size_t* bytes) = 0;
  virtual int sum_partials_device(const void* d_words, uint64_t n, uint8_t* out, int* is_inf) = 0;
  virtual int poly_quotient(const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* q_out) = 0;
  virtual int poly_eval(const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* y_out) = 0;
  virtual int prove(const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* out, int* is_inf) = 0;
  virtual int prove_batch(const uint8_t* coeffs, uint64_t n, const int64_t* first_roots, uint64_t k, uint64_t batch,
                          uint8_t* out, int* is_inf) = 0;
  virtual int interpolate(const uint8_t* ys, int64_t first_x, uint64_t k, uint8_t* coeffs_out, uint8_t* vanish_out) = 0;
  virtual int commit_bytes(const uint8_t* bytes, uint64_t byte_length, int chunk_size, int64_t chunk_offset,
                           uint8_t* coeffs_out, uint8_t* out, int* is_inf) = 0;
  virtual int microbench(int kind, double* ops) = 0;
  virtual int selftest(int* failures) = 0;
  virtual int phase_times(float* ms, int n) = 0;
  virtual int msm_stats(uint64_t* out, int n) = 0;

  // Cooperative launches (msm_tail.cuh): every CTA must be resident, so the grid is capped at what the occupancy
  // calculator allows for that kernel on this device (cached per kernel).
  std::map<const void*, int> coop_blocks;
  int num_sms = 0;
  int max_cooperative_blocks(const void* kernel, int threads) {
    auto it = coop_blocks.find(kernel);
    if (it != coop_blocks.end()) return it->second;
    if (!num_sms && cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) num_sms = 0;
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess) per_sm = 0;
    const int v = per_sm * num_sms;
    coop_blocks[kernel] = v;
    return v;
  }
  int launch_cooperative(const void* kernel, uint64_t blocks_wanted, int threads, void* args_struct, size_t smem = 0,
                         void* second_arg = nullptr) {
    int cap = max_cooperative_blocks(kernel, threads);
    if (cap <= 0) { last_error = "cooperative launch: kernel does not fit the device"; return KZG_GPU_ERR_CUDA; }
    if (num_sms > 0) cap = std::min(cap, num_sms);  // one CTA per SM: cheap grid barriers, one warp per sub-partition
    const unsigned grid = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(blocks_wanted, (uint64_t)cap));
    void* params[2] = {args_struct, second_arg};
    cudaError_t e = cudaLaunchCooperativeKernel(kernel, dim3(grid), dim3((unsigned)threads), params, smem, stream);
    if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchCooperativeKernel");
    launches++;
    return KZG_GPU_OK;
  }

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
  DevBuf g1_table, g2_table; // per-SRS window tables T[w][i] = 2^(c*w) * G_i (row 0 = the SRS itself)
  int table_c = 0, table_W = 0;
  size_t g1_point_stride = sizeof(Affine<Fp>);
  // The table ladder: window tables of PREFIXES of a large SRS with the window width that suits the prefix.  The window
  // width of the main table is chosen for srs_n (c = 20: 2^19 buckets at 2^20 points), and an MSM walks the whole bucket
  // reduction of that width however few terms it has — a degree-4096 commitment against a 2^20-point setup (the
  // reference's benchmark.cpp commits polynomials of degree 1024 and up against one 10.4 M-term setup) would cost
  // ~1.2 ms instead of the 0.46 ms it costs against a 4096-point one.  MSMs over points [0, n_k) use rung k.
  struct TableRung { DevBuf buf; uint64_t n = 0; int c = 0, W = 0; };
  static constexpr int N_RUNGS = 2;
  static constexpr uint64_t RUNG_POINTS[N_RUNGS] = {8192, 131072};
  TableRung rungs[N_RUNGS];
  // Direct-lookup table (msm_direct.cuh): D[w][i][d-1] = d 2^(c w) G_i for every digit value, for a SMALL SRS when the
  // caller grants the memory ("direct_table_gb").  G1 MSMs then need no sort, no buckets and no weighted reduction.
  DevBuf g1_direct;
  int direct_c = 0, direct_W = 0;
  // fixed-base tables of the generators
  // (constants of the curve: built once per process and device, shared by every context, never freed)
  struct FbShared { Affine<Fp>* t1 = nullptr; Affine<Fp2>* t2 = nullptr; };
  static std::mutex& fb_mutex() { static std::mutex m; return m; }
  static std::map<int, FbShared>& fb_cache() { static std::map<int, FbShared> c; return c; }
  Affine<Fp>* fb1 = nullptr;
  Affine<Fp2>* fb2 = nullptr;
  DevBuf fb_sums;
  bool fb_ready = false;
  // scratch
  DevBuf nt_g0, nt_g1, nt_u, raw, stage, scal2, scal, digits, counts, starts, task_first, cursor, sorted, partial, group_out, out_bytes,
      out_inf, out_xyzz, poly0, poly1, qbytes, powers, tile_sums, hist, tasks, slot_seg, ranks, red0, red1, red2, red3, aff_off, aff_a, aff_b, aff_prefix,
      aff_totals, roots, inv_fact, it_u, it_pa, it_pb, it_za, it_zb;
  uint64_t inv_fact_n = 0;
  // description of the most recent launch, for fetch()
  uint64_t last_batch = 0;
  int last_group = 1;

  ~Ctx() override {
    cudaSetDevice(device);
    if (aux_stream) { cudaStreamSynchronize(aux_stream); cudaStreamDestroy(aux_stream); }
    if (g2_table_ev) cudaEventDestroy(g2_table_ev);
    for (DevBuf* b : {&barrier_buf, &g1, &g2, &g1_table, &g2_table, &fb_sums, &nt_g0, &nt_g1, &nt_u, &raw, &stage, &scal2, &scal, &digits, &counts, &starts, &task_first,
                      &cursor, &sorted, &partial, &group_out, &out_bytes, &out_inf, &out_xyzz, &poly0, &poly1,
                      &qbytes, &powers, &tile_sums, &hist, &tasks, &slot_seg, &ranks, &aff_off, &aff_a, &aff_b, &aff_prefix, &aff_totals, &roots, &inv_fact, &it_u, &it_pa, &it_pb, &it_za, &it_zb, &red0, &red1, &red2, &red3})
      b->release();
    for (cudaEvent_t e : phase_ev) if (e) cudaEventDestroy(e);
    for (int i = 0; i < 2; i++) {
      if (copied_ev[i]) cudaEventDestroy(copied_ev[i]);
      if (consumed_ev[i]) cudaEventDestroy(consumed_ev[i]);
    }
    for (cudaEvent_t e : chunk_ev) if (e) cudaEventDestroy(e);
    if (copy_stream) cudaStreamDestroy(copy_stream);
    if (stream) cudaStreamDestroy(stream);
  }

  int init() {
    CK(cudaSetDevice(device));
    CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    for (cudaEvent_t& e : phase_ev) CK(cudaEventCreate(&e));
    return KZG_GPU_OK;
  }

  // ------------------------------------------------------------------ SRS
  // window width of the per-SRS table: minimise  n*W*10 + 2^(c-1)*30  modmuls (mixed adds + bucket
  // reduction) for an MSM over the whole SRS, subject to W <= MAXW and a table-size budget.  A width whose
  // LAST window holds only a few scalar bits sends n / 2^(bits-1) records to each of a handful of buckets:
  // their histogram atomics serialise (measured 0.27 ms for 2048 records per bucket at n = 2^14, c = 14) and
  // their task lists need deep folds, so that pile-up is charged at ~3800 modmul-equivalents per record.
  int pick_table_c(uint64_t n) const {
    if (opt_window_bits) return opt_window_bits;
    int best = 0;
    double best_cost = 0;
    for (int c = 7; c <= 21; c++) {
      const int W = (C::SCALAR_BITS + 1 + c - 1) / c;
      if (W > MAXW) continue;
      if ((double)W * n * sizeof(Affine<Fp>) > 24e9) continue;
      const int top_bits = C::SCALAR_BITS + 1 - (W - 1) * c;  // >= 1
      const double pile = (double)n / (double)(1ull << (top_bits - 1));
      const double cost = (double)n * W * 10.0 + (double)(1ull << (c - 1)) * 30.0 + 3800.0 * pile;
      if (!best || cost < best_cost) { best = c; best_cost = cost; }
    }
    return best ? best : 8;
  }
  // window width without the table (every window has its own bucket group and the finish does the doublings):
  // the same cost model with W bucket groups to reduce
  int pick_plain_c(uint64_t count) const {
    if (count < 32) return 4;
    int best = 8;
    double best_cost = 0;
    for (int c = 6; c <= 16; c++) {
      const int W = (C::SCALAR_BITS + 1 + c - 1) / c;
      const int top_bits = C::SCALAR_BITS + 1 - (W - 1) * c;
      const double pile = (double)count / (double)(1ull << (top_bits - 1));
      const double cost = (double)count * W * 10.0 + (double)W * (double)(1ull << (c - 1)) * 30.0 + 3800.0 * pile;
      if (best_cost == 0 || cost < best_cost) { best = c; best_cost = cost; }
    }
    return best;
  }
  int build_table() {
    if (aux_stream) CK(cudaStreamSynchronize(aux_stream));  // a table build of the previous SRS may still be running
    g2_table_pending = false;
    g1_table.release();
    g2_table.release();
    for (TableRung& rg : rungs) { rg.buf.release(); rg.n = 0; }
    g1_direct.release();
    direct_c = direct_W = 0;
    table_c = table_W = 0;
    if (srs_n == 0) return KZG_GPU_OK;
    if (!opt_precompute) return reserve_scratch();
    const int c = pick_table_c(srs_n);
    const int W = (C::SCALAR_BITS + 1 + c - 1) / c;
    if (W > MAXW) {
      last_error = "window_bits is too small for the per-SRS window table (more than 40 windows); use precompute = 0 with it";
      return KZG_GPU_ERR_ARG;
    }
    // "table_pad" experiment: one 128-byte line per 96-byte point (BLS12-381), so a gathered point never straddles two lines
    g1_point_stride = opt_table_pad && sizeof(Affine<Fp>) <= 128 ? 128 : sizeof(Affine<Fp>);
    CK(g1_table.ensure((size_t)W * srs_n * g1_point_stride));
    if (srs_n <= TABLE_TEAM_MAX_POINTS)  // latency-bound build: four lanes per point (setup.cuh)
      msm_precompute_table_team<Fp, MAXW><<<(unsigned)((TeamWidth<Fp>::LANES * srs_n + 127) / 128), 128, 0, stream>>>(
          srs_n, srs_n, c, W, g1.as<Affine<Fp>>(), g1_table.as<Affine<Fp>>(), g1_point_stride);
    else
      msm_precompute_table<Fp, MAXW><<<(unsigned)((srs_n + 127) / 128), 128, 0, stream>>>(
          srs_n, srs_n, c, W, g1.as<Affine<Fp>>(), g1_table.as<Affine<Fp>>(), g1_point_stride);
    LAUNCHED();
    for (int k = 0; k < N_RUNGS; k++) {
      TableRung& rg = rungs[k];
      rg.buf.release();
      rg.n = 0;
      const uint64_t nk = RUNG_POINTS[k];
      if (!opt_table_ladder || opt_window_bits || srs_n < 4 * nk) continue;
      const int ck = pick_table_c(nk), Wk = (C::SCALAR_BITS + 1 + ck - 1) / ck;
      if (ck >= c || Wk > MAXW) continue;
      CK(rg.buf.ensure((size_t)Wk * nk * sizeof(Affine<Fp>)));
      if (nk <= TABLE_TEAM_MAX_POINTS)
        msm_precompute_table_team<Fp, MAXW><<<(unsigned)((TeamWidth<Fp>::LANES * nk + 127) / 128), 128, 0, stream>>>(
            nk, nk, ck, Wk, g1.as<Affine<Fp>>(), rg.buf.template as<Affine<Fp>>(), sizeof(Affine<Fp>));
      else
        msm_precompute_table<Fp, MAXW><<<(unsigned)((nk + 127) / 128), 128, 0, stream>>>(
            nk, nk, ck, Wk, g1.as<Affine<Fp>>(), rg.buf.template as<Affine<Fp>>(), sizeof(Affine<Fp>));
      LAUNCHED();
      rg.n = nk; rg.c = ck; rg.W = Wk;
    }
    if (opt_direct_table_gb > 0 && !opt_window_bits) {
      // the widest window (fewest additions per scalar) whose table fits the memory the caller granted
      const double budget = (double)opt_direct_table_gb * 1073741824.0;
      int cd = 0, Wd = 0;
      for (int cc = 16; cc >= 8; cc--) {
        const int Wc = (C::SCALAR_BITS + 1 + cc - 1) / cc;
        if (Wc > MAXW) continue;
        if ((double)Wc * (double)srs_n * (double)(1ull << (cc - 1)) * sizeof(Affine<Fp>) <= budget) { cd = cc; Wd = Wc; break; }
      }
      if (cd) {  // the grant is an upper bound, not a promise: with the memory taken (other resident setups) the SRS simply has no such table
        const cudaError_t ed = g1_direct.ensure_exact((size_t)Wd * srs_n * (1ull << (cd - 1)) * sizeof(Affine<Fp>));
        if (ed == cudaErrorMemoryAllocation) { cudaGetLastError(); cd = 0; }
        else if (ed != cudaSuccess) return cuda_fail(ed, "g1_direct.ensure");
      }
      if (cd) {
        const uint64_t H = 1ull << (cd - 1), rows = (uint64_t)Wd * srs_n;
        // entry d = 1 of every row: 2^(cd w) G_i, written by the window-table kernel with one row of H entries as its stride
        if (srs_n <= TABLE_TEAM_MAX_POINTS)
          msm_precompute_table_team<Fp, MAXW><<<(unsigned)((TeamWidth<Fp>::LANES * srs_n + 127) / 128), 128, 0, stream>>>(
              srs_n, srs_n, cd, Wd, g1.as<Affine<Fp>>(), g1_direct.as<Affine<Fp>>(), (size_t)H * sizeof(Affine<Fp>));
        else
          msm_precompute_table<Fp, MAXW><<<(unsigned)((srs_n + 127) / 128), 128, 0, stream>>>(
              srs_n, srs_n, cd, Wd, g1.as<Affine<Fp>>(), g1_direct.as<Affine<Fp>>(), (size_t)H * sizeof(Affine<Fp>));
        LAUNCHED();
        for (int level = 0; level + 1 < cd; level++) {
          const uint64_t total = rows << level;
          direct_table_fill<Fp><<<(unsigned)std::min<uint64_t>((total + 127) / 128, (uint64_t)148 * 256), 128, 0, stream>>>(
              level, rows, H, g1_direct.as<Affine<Fp>>());
          LAUNCHED();
        }
        direct_c = cd;
        direct_W = Wd;
      }
    }
    CK(cudaStreamSynchronize(stream));
    table_c = c;
    table_W = W;
    return reserve_scratch();
  }
  // The G2 window table is built LAZILY, by the first G2 MSM of at least G2_TABLE_MIN_TERMS terms: G2 MSMs only carry the
  // k + 1 coefficients of a vanishing polynomial (verify_proof, src/trusted_setup.cpp:241), a setup that is never
  // asked for a many-point verification never pays for the table (it was most of trusted_setup(4096) through the
  // drop-in in round 1: 36 ms on BLS12-381), and it is not built at all for SRS sizes where it would take gigabytes.
  // For the SRS sizes the reference's callers use (<= 65536 points) the build starts when the SRS becomes resident, on a
  // SIDE stream, so that neither the constructor nor the commitments that follow wait for it; the first G2 MSM does.
  static constexpr uint64_t TABLE_TEAM_MAX_POINTS = 16384;  // up to ~one team per warp slot of the machine
  static constexpr uint64_t G2_TABLE_MIN_TERMS = 256;
  static constexpr uint64_t G2_TABLE_EAGER_MAX_SRS = 65536;
  cudaStream_t aux_stream = nullptr;
  cudaEvent_t g2_table_ev = nullptr;
  bool g2_table_pending = false;
  int ensure_g2_table(uint64_t count, bool eager = false) {
    if (g2_table.p || !have_g2 || opt_precompute < 2 || !table_c) return KZG_GPU_OK;
    if (!eager && count < G2_TABLE_MIN_TERMS) return KZG_GPU_OK;
    if (eager && srs_n > G2_TABLE_EAGER_MAX_SRS) return KZG_GPU_OK;
    if ((double)table_W * (double)srs_n * sizeof(Affine<Fp2>) > 8e9) return KZG_GPU_OK;
    CK(g2_table.ensure((size_t)table_W * srs_n * sizeof(Affine<Fp2>)));
    if (!aux_stream) CK(cudaStreamCreateWithFlags(&aux_stream, cudaStreamNonBlocking));
    if (!g2_table_ev) CK(cudaEventCreateWithFlags(&g2_table_ev, cudaEventDisableTiming));
    CK(cudaEventRecord(g2_table_ev, stream));          // the SRS itself was written on the main stream
    CK(cudaStreamWaitEvent(aux_stream, g2_table_ev, 0));
    if (srs_n <= TABLE_TEAM_MAX_POINTS)
      msm_precompute_table_team<Fp2, MAXW><<<(unsigned)((TeamWidth<Fp2>::LANES * srs_n + 127) / 128), 128, 0, aux_stream>>>(
          srs_n, srs_n, table_c, table_W, g2.as<Affine<Fp2>>(), g2_table.as<Affine<Fp2>>(), sizeof(Affine<Fp2>));
    else
      msm_precompute_table<Fp2, MAXW><<<(unsigned)((srs_n + 63) / 64), 64, 0, aux_stream>>>(
          srs_n, srs_n, table_c, table_W, g2.as<Affine<Fp2>>(), g2_table.as<Affine<Fp2>>(), sizeof(Affine<Fp2>));
    LAUNCHED();
    CK(cudaEventRecord(g2_table_ev, aux_stream));
    g2_table_pending = true;
    return KZG_GPU_OK;
  }
  int wait_g2_table() {  // before the main stream reads (or frees) the table
    if (g2_table_pending) {
      CK(cudaStreamWaitEvent(stream, g2_table_ev, 0));
      g2_table_pending = false;
    }
    return KZG_GPU_OK;
  }

  // Size the MSM scratch for one MSM over the whole SRS now, so that the first commitment of each new size does
  // not pay ~20 cudaFree/cudaMalloc pairs (measured: 1.9 ms instead of 1.0 ms for the first degree-4096 commit
  // after a degree-2048 one in the reference's benchmark.cpp).  Larger batches still grow the buffers on demand.
  // A very large SRS reserves for MSMs of up to 2^20 terms (0.3 GB): the reference's benchmark.cpp commits
  // polynomials of degree 1024 .. 2^23 against one 10.4 M-term setup, and growing twenty buffers of tens of MB in front
  // of each new size cost 10-20 ms per row (profiles/r2common_benchmark_common_bls12381.txt).
  int reserve_scratch() {
    if (srs_n == 0) return KZG_GPU_OK;
    const uint64_t cnt = std::min<uint64_t>(srs_n, srs_n > (1ull << 22) ? (1ull << 20) : srs_n);
    CK(scal.ensure(cnt * (size_t)SB));
    int rc = enqueue_msm<Fp>(g1.as<Affine<Fp>>(), g1_table.p ? g1_table.as<Affine<Fp>>() : nullptr, table_c, nullptr, cnt, 1, 0, true);
    if (rc) return rc;
    if (have_g2 && srs_n <= (1ull << 16))  // verify-side MSMs (k + 1 terms)
      rc = enqueue_msm<Fp2>(g2.as<Affine<Fp2>>(), table_c ? g2.as<Affine<Fp2>>() : nullptr, table_c, nullptr, cnt, 1, 0, true);
    if (rc) return rc;
    if (g1_direct.p) {  // the partial-sum slots of one direct-lookup MSM over the whole SRS (msm_direct.cuh)
      const DirectShape ds = direct_shape(cnt, 1, 0);
      CK(partial.ensure(ds.P2 * sizeof(XYZZ<Fp>)));
      CK(red0.ensure(std::max<uint64_t>(1, ds.P2 / 2) * sizeof(XYZZ<Fp>)));
      CK(red2.ensure(sizeof(XYZZ<Fp>)));
      CK(red3.ensure(sizeof(XYZZ<Fp>)));
    }
    return warm_up();
  }
  // One two-term MSM of zero scalars per group right after the SRS became resident: the first launch of every kernel of
  // the pipeline loads its code (CUDA loads modules lazily) and sets its attributes here instead of inside the first
  // create_commit (round 1: the first commitment of a process took 1.0 ms against 0.5 ms for the second).  Then the
  // G2 window table starts building on the side stream, and the 1/t! table of the interpolation on the main one.
  int warm_up() {
    const uint64_t cnt = std::min<uint64_t>(2, srs_n);
    CK(scal.ensure(cnt * (size_t)SB));
    CK(cudaMemsetAsync(scal.p, 0, cnt * (size_t)SB, stream));
    int rc = enqueue_msm<Fp>(g1.as<Affine<Fp>>(), g1_table.p ? g1_table.as<Affine<Fp>>() : nullptr, table_c, scal.as<uint8_t>(), cnt, 1, 0);
    if (!rc && have_g2) rc = enqueue_msm<Fp2>(g2.as<Affine<Fp2>>(), nullptr, table_c, scal.as<uint8_t>(), cnt, 1, 0);
    if (rc) return rc;
    last_batch = 0;
    if (inv_fact_n < 4096 && srs_n <= G2_TABLE_EAGER_MAX_SRS) {
      CK(inv_fact.ensure(4096 * sizeof(Fr)));
      fr_inv_factorials<C><<<1, INV_FACT_THREADS, 0, stream>>>(4096u, inv_fact.as<Fr>());
      LAUNCHED();
      inv_fact_n = 4096;
    }
    return ensure_g2_table(0, /*eager=*/true);
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

  // ------------------------------------------------------------------ MSM
  // exclusive scan of counts (or of ceil(counts/chunk)) into out[0..n], three launches
  int enqueue_scan(uint64_t n, int chunk, const uint32_t* in, uint32_t* out, uint32_t* zero_me) {
    const uint64_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    CK(tile_sums.ensure((ntiles + 1) * 4));
    scan_tile_sums<<<(unsigned)ntiles, SCAN_THREADS, 0, stream>>>(n, chunk, in, tile_sums.as<uint32_t>());
    LAUNCHED();
    scan_tiles<<<1, 1024, 0, stream>>>(ntiles, tile_sums.as<uint32_t>());
    LAUNCHED();
    scan_apply<<<(unsigned)ntiles, SCAN_THREADS, 0, stream>>>(n, chunk, in, tile_sums.as<uint32_t>(), out, zero_me);
    LAUNCHED();
    return KZG_GPU_OK;
  }

  // pairwise tree levels over two independent arrays (rows / columns side by side): every sum holds
  // len[k] contiguous values; halves them until one value per sum is left.
  template <class F>
  int enqueue_pair_trees(XYZZ<F>*& in0, XYZZ<F>*& out0, uint64_t sums0, uint64_t len0, XYZZ<F>*& in1, XYZZ<F>*& out1,
                         uint64_t sums1, uint64_t len1) {
    while (len0 > 1 || len1 > 1) {
      PairJob<F> j0{in0, out0, len0 > 1 ? sums0 * (len0 / 2) : 0};
      PairJob<F> j1{in1, out1, len1 > 1 ? sums1 * (len1 / 2) : 0};
      const uint64_t mx = std::max(j0.n_out, j1.n_out);
      if (opt_team && j0.n_out + j1.n_out <= TEAM_MAX_ITEMS) {  // latency-bound level: 4 lanes per addition
        dim3 grid((unsigned)((WIDE_LANES * mx + 127) / 128), 2);
        msm_pair_sum_team<F><<<grid, 128, 0, stream>>>(j0, j1);
      } else {
        dim3 grid((unsigned)std::min<uint64_t>((mx + 127) / 128, (uint64_t)148 * 64), 2);
        msm_pair_sum<F><<<grid, 128, 0, stream>>>(j0, j1);
      }
      LAUNCHED();
      if (len0 > 1) { std::swap(in0, out0); len0 /= 2; }
      if (len1 > 1) { std::swap(in1, out1); len1 /= 2; }
    }
    return KZG_GPU_OK;
  }

  // msm_tail<F>: one CTA of TAIL_THREADS per SM; the dynamic shared-memory request only keeps a second CTA off the SM
  template <class F>
  int prepare_tail() {
    if (tail_ready[std::is_same<F, Fp2>::value ? 1 : 0]) return ensure_barrier();
    CK(cudaFuncSetAttribute(msm_tail<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, TAIL_SMEM_BYTES));
    CK(cudaFuncSetAttribute(msm_tail_proj<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, TAIL_SMEM_BYTES));
    int rcb = ensure_barrier();
    if (rcb) return rcb;
    tail_ready[std::is_same<F, Fp2>::value ? 1 : 0] = true;
    return KZG_GPU_OK;
  }
  int ensure_barrier() {
    if (!barrier_buf.p) {
      CK(barrier_buf.ensure(sizeof(GridBarrier)));
      CK(cudaMemsetAsync(barrier_buf.p, 0, sizeof(GridBarrier), stream));
      barrier_sleep_set = -1;
    }
    if (barrier_sleep_set != opt_barrier_sleep) {
      const unsigned v = (unsigned)opt_barrier_sleep;
      CK(cudaMemcpyAsync(reinterpret_cast<char*>(barrier_buf.p) + offsetof(GridBarrier, sleep_ns), &v, sizeof v,
                         cudaMemcpyHostToDevice, stream));
      CK(cudaStreamSynchronize(stream));  // `v` is a stack variable
      barrier_sleep_set = opt_barrier_sleep;
    }
    return KZG_GPU_OK;
  }
  int barrier_sleep_set = -1;
  bool tail_ready[2] = {false, false};
  DevBuf barrier_buf;
  uint64_t tail_threads_total() {
    if (!num_sms) cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, device);
    return (uint64_t)(num_sms > 0 ? num_sms : 1) * TAIL_THREADS;
  }

  void phase_mark(int i) {
    if (opt_profile && phase_ev[0]) cudaEventRecord(phase_ev[i], stream);
  }

  // Enqueue one Pippenger pass for `batch` MSMs of `count` scalars (device-resident bytes).
  // Leaves affine bytes + flags in out_bytes/out_inf and the projective result of each MSM in out_xyzz.
  template <class F>
  // table_rows: points per row of `table` (0: srs_n — the main table; otherwise a rung of the table ladder)
  int enqueue_msm(const Affine<F>* bases, const Affine<F>* table, int tc, const uint8_t* d_scalars, uint64_t count,
                  uint64_t batch, uint64_t base_offset, bool reserve_only = false, uint64_t table_rows = 0) {
    MsmShape sh;
    const bool precomp = table != nullptr;
    sh.c = precomp ? tc : (opt_window_bits ? opt_window_bits : pick_plain_c(count));
    sh.W = (C::SCALAR_BITS + 1 + sh.c - 1) / sh.c;
    sh.nb = 1 << (sh.c - 1);
    sh.precomp = precomp ? 1 : 0;
    sh.groups_per_msm = precomp ? 1 : sh.W;
    sh.n = count;
    sh.batch = batch;
    sh.table_stride = table_rows ? table_rows : srs_n;
    sh.base_offset = base_offset;
    const uint64_t records = count * batch * (uint64_t)sh.W;
    // task length: long enough to amortise the fold, short enough that a small MSM still fills the machine
    // Batched-affine pre-reduction levels ("affine_levels" option).  OFF by default: measured on B200 at 2^20
    // (profiles/r1f_affine_levels.md) two levels cost 6.1 ms against the 3.7 ms of XYZZ additions they replace —
    // each level gathers its inputs twice (forward and backward pass) from the 1.25 GiB window table and is bound
    // by random-gather bandwidth, while the XYZZ kernel hides one gather behind 10 modmuls.
    int levels = 0;
    if (std::is_same<F, Fp>::value) levels = opt_affine_levels >= 0 ? opt_affine_levels : 0;
    if (g1_point_stride != sizeof(Affine<Fp>)) levels = 0;  // the affine levels read the table at its natural stride
    if (table_rows) levels = 0;
    sh.chunk = opt_task_chunk ? std::min(opt_task_chunk, MSM_MAX_CHUNK)
                              : (int)std::min<uint64_t>(128, std::max<uint64_t>(4, (records >> levels) / 65536));
    if (records >= (1ull << 31) || (precomp && (uint64_t)sh.W * sh.table_stride >= (1ull << 31))) return KZG_GPU_ERR_ARG;
    const uint64_t nbuckets = sh.total_buckets();
    const uint64_t max_tasks = ((records >> levels) + nbuckets) / sh.chunk + nbuckets + 1;
    const uint64_t groups = sh.total_groups();
    const int h = (sh.c - 1) / 2;
    const uint64_t Cn = 1ull << h, R = (uint64_t)sh.nb / Cn;
    const uint64_t red_cap = groups * ((uint64_t)sh.nb / 2 + R + Cn + 2);

    CK(digits.ensure(records * 4));
    CK(ranks.ensure(records * 4));
    CK(sorted.ensure(records * 4));
    CK(counts.ensure((nbuckets + 1) * 4));
    CK(starts.ensure((nbuckets + 1) * 4));
    CK(task_first.ensure((nbuckets + 1) * 4));
    if (levels > 0) {
      CK(aff_off.ensure((size_t)levels * (nbuckets + 1) * 4));
      CK(aff_a.ensure(((records >> 1) + nbuckets) * sizeof(Affine<F>)));
      if (levels > 1) CK(aff_b.ensure(((records >> 2) + nbuckets) * sizeof(Affine<F>)));
      CK(aff_prefix.ensure(((records >> 1) + nbuckets) * sizeof(F)));
      CK(aff_totals.ensure((((records >> 1) + nbuckets) / AFF_K + 2) * sizeof(F)));
    }
    CK(hist.ensure((3 * (MSM_MAX_CHUNK + 1) + 1) * 4));
    CK(slot_seg.ensure(max_tasks * sizeof(MsmSlotSeg)));
    CK(tasks.ensure(max_tasks * sizeof(MsmTask)));
    CK(partial.ensure(max_tasks * sizeof(XYZZ<F>)));
    for (DevBuf* rb : {&red0, &red1, &red2, &red3}) CK(rb->ensure(red_cap * sizeof(XYZZ<F>)));
    CK(group_out.ensure(groups * sizeof(XYZZ<F>)));
    CK(out_bytes.ensure(batch * (size_t)(4 * M)));
    CK(out_inf.ensure(batch * sizeof(int)));
    CK(out_xyzz.ensure(batch * sizeof(XYZZ<F>)));
    if (reserve_only) return KZG_GPU_OK;  // srs_load / setup_powers size the scratch for a full-SRS MSM up front
    uint32_t* d_hist = hist.as<uint32_t>();
    uint32_t* d_hist_start = d_hist + (MSM_MAX_CHUNK + 2);  // [MSM_MAX_CHUNK + 1] = most tasks in one bucket
    uint32_t* d_hist_cursor = d_hist_start + (MSM_MAX_CHUNK + 1);

    const int sms = 148;
    // Small and medium MSMs (and the narrow end of every large one) run their many short steps inside two
    // persistent cooperative kernels (msm_tail.cuh) instead of ~50 launches: "tail" option, on by default.
    const bool use_tail = opt_tail && opt_team;
    if (use_tail) {
      int rc0 = prepare_tail<F>();
      if (rc0) return rc0;
    }
    const bool fused_sort = use_tail && chunk_slices <= 1 && levels == 0 && records <= (1ull << 22) && nbuckets <= (1ull << 22);
    phase_mark(0);
    int rc = KZG_GPU_OK;
    const uint32_t* seg_off = starts.as<uint32_t>();
    const uint32_t* seg_sorted = sorted.as<uint32_t>();
    const Affine<F>* seg_pts = precomp ? table : bases;
    if (fused_sort) {
      chunk_slices = 0;
      const uint64_t ntiles = (nbuckets + SCAN_TILE - 1) / SCAN_TILE;
      CK(tile_sums.ensure((ntiles + 1) * 4));
      SortArgs sa{sh, barrier_buf.as<GridBarrier>(), d_scalars, digits.as<int32_t>(), ranks.as<uint32_t>(), counts.as<uint32_t>(), starts.as<uint32_t>(),
                  task_first.as<uint32_t>(), sorted.as<uint32_t>(), tile_sums.as<uint32_t>(), d_hist, d_hist_start,
                  d_hist_cursor, tasks.as<MsmTask>(), slot_seg.as<MsmSlotSeg>(), nbuckets, sh.chunk};
      const uint64_t want = (std::max<uint64_t>(records, nbuckets) + SCAN_THREADS - 1) / SCAN_THREADS;
      rc = launch_cooperative((const void*)msm_sort_small<C>, want, SCAN_THREADS, &sa);  // at most one CTA per SM
      if (rc) return rc;
    } else {
    CK(cudaMemsetAsync(counts.p, 0, (nbuckets + 1) * 4, stream));
    CK(cudaMemsetAsync(hist.p, 0, (MSM_MAX_CHUNK + 2) * 4, stream));
    {
      const uint64_t total = count * batch;
      const int slices = chunk_slices > 1 ? chunk_slices : 1;  // > 1: the scalars arrive in that many H2D chunks
      for (int sl = 0; sl < slices; sl++) {
        const uint64_t t0 = total * sl / slices, t1 = total * (sl + 1) / slices;
        if (slices > 1) CK(cudaStreamWaitEvent(stream, chunk_ev[sl], 0));
        unsigned grid = (unsigned)std::min<uint64_t>((t1 - t0 + 255) / 256, (uint64_t)sms * 16);
        msm_digits<C><<<grid, 256, 0, stream>>>(sh, d_scalars, digits.as<int32_t>(), ranks.as<uint32_t>(), counts.as<uint32_t>(),
                                                t0, t1);
        LAUNCHED();
      }
      chunk_slices = 0;
    }
    rc = enqueue_scan(nbuckets, 0, counts.as<uint32_t>(), starts.as<uint32_t>(), nullptr);
    if (rc) return rc;
    {
      unsigned grid = (unsigned)std::min<uint64_t>((records + 255) / 256, (uint64_t)sms * 32);
      msm_scatter<<<grid, 256, 0, stream>>>(sh, digits.as<int32_t>(), starts.as<uint32_t>(), ranks.as<uint32_t>(),
                                            sorted.as<uint32_t>());
      LAUNCHED();
    }
    // batched-affine pre-reduction (G1 only): `levels` rounds of pairwise affine additions inside the buckets
    if constexpr (std::is_same<F, Fp>::value) {
      for (int l = 1; l <= levels; l++) {
        uint32_t* off_l = aff_off.as<uint32_t>() + (size_t)(l - 1) * (nbuckets + 1);
        rc = enqueue_scan(nbuckets, 1 << l, counts.as<uint32_t>(), off_l, nullptr);
        if (rc) return rc;
        const uint64_t cap = (records >> l) + nbuckets;  // bound on sum ceil(cnt / 2^l)
        const uint64_t nthreads = (cap + AFF_K - 1) / AFF_K;
        AffLevel<F> lv{seg_sorted, seg_pts, seg_off, off_l, nbuckets, (l & 1) ? aff_a.as<Affine<F>>() : aff_b.as<Affine<F>>(),
                       aff_prefix.as<F>(), aff_totals.as<F>()};
        aff_forward<F><<<(unsigned)((nthreads + 127) / 128), 128, 0, stream>>>(lv, nthreads);
        LAUNCHED();
        constexpr int K2 = 4;
        const uint64_t inv_threads = (nthreads + K2 - 1) / K2;
        aff_invert<F, K2><<<(unsigned)((inv_threads + 127) / 128), 128, 0, stream>>>(off_l + nbuckets, aff_totals.as<F>());
        LAUNCHED();
        aff_backward<F><<<(unsigned)((nthreads + 127) / 128), 128, 0, stream>>>(lv, nthreads);
        LAUNCHED();
        seg_sorted = nullptr;
        seg_pts = lv.out_pts;
        seg_off = off_l;
      }
    }
    rc = enqueue_scan(nbuckets, sh.chunk << levels, counts.as<uint32_t>(), task_first.as<uint32_t>(), nullptr);
    if (rc) return rc;
    {
      unsigned grid = (unsigned)std::min<uint64_t>((nbuckets + 255) / 256, (uint64_t)sms * 8);
      msm_task_hist<<<grid, 256, 0, stream>>>(nbuckets, sh.chunk, seg_off, task_first.as<uint32_t>(), d_hist);
      LAUNCHED();
      msm_task_hist_scan<<<1, 32, 0, stream>>>(d_hist, d_hist_start, d_hist_cursor);
      LAUNCHED();
      msm_task_place<<<grid, 256, 0, stream>>>(nbuckets, sh.chunk, seg_off, task_first.as<uint32_t>(),
                                               d_hist_start, d_hist_cursor, tasks.as<MsmTask>(),
                                               slot_seg.as<MsmSlotSeg>());
      LAUNCHED();
    }
    }
    phase_mark(1);
    {
      unsigned grid = (unsigned)std::min<uint64_t>((max_tasks + 127) / 128, (uint64_t)1 << 30);
      const size_t pt_stride = (precomp && std::is_same<F, Fp>::value && levels == 0 && !table_rows) ? g1_point_stride : sizeof(Affine<F>);
      msm_accumulate<F><<<grid, 128, 0, stream>>>(task_first.as<uint32_t>() + nbuckets, tasks.as<MsmTask>(), seg_sorted,
                                                  seg_pts, partial.as<XYZZ<F>>(), pt_stride);
      LAUNCHED();
    }
    phase_mark(2);
    XYZZ<F>*rin = red0.as<XYZZ<F>>(), *rout = red1.as<XYZZ<F>>(), *cin = red2.as<XYZZ<F>>(), *cout = red3.as<XYZZ<F>>();
    const uint64_t Lr = Cn >= 2 ? Cn / 2 : 1, Lc = R >= 2 ? R / 2 : 1;
    const uint64_t Pr = R >= WCHUNK ? R / WCHUNK : 1, Pc = Cn >= WCHUNK ? Cn / WCHUNK : 1;
    // The reduction is a fixed sequence of steps whose item counts only shrink.  Steps wider than the tail
    // kernel's team count are machine-filling launches of their own; from the first step that fits, everything
    // that is left runs inside ONE cooperative kernel (msm_tail) with grid barriers between the steps.
    // a step of up to four rounds of one group operation per thread (msm_tail.cuh) belongs in the tail
    const uint64_t tail_cap = use_tail ? 4 * tail_threads_total() : 0;
    const bool finish_in_tail = use_tail && batch <= tail_threads_total() / 32;  // one MSM per warp, one round (see msm_tail)
    TailArgs<F> ta{};
    ta.barrier = barrier_buf.as<GridBarrier>();
    ta.ntasks_ptr = task_first.as<uint32_t>() + nbuckets;
    ta.max_nt = d_hist + MSM_MAX_CHUNK + 1;
    ta.slot_seg = slot_seg.as<MsmSlotSeg>();
    ta.partial = partial.as<XYZZ<F>>();
    ta.task_first = task_first.as<uint32_t>();
    ta.nb = (uint64_t)sh.nb;
    ta.groups = groups; ta.R = R; ta.Cn = Cn; ta.h = h;
    ta.group_out = group_out.as<XYZZ<F>>();
    ta.batch = batch; ta.groups_per_msm = sh.groups_per_msm; ta.c = sh.c; ta.modbytes = M; ta.want_affine = opt_finish_affine;
    ta.out_bytes = out_bytes.as<uint8_t>(); ta.is_inf = out_inf.as<int>(); ta.out_xyzz = out_xyzz.as<XYZZ<F>>();
    auto launch_tail = [&](int first, int last, uint64_t r_len, uint64_t c_len, uint64_t items) -> int {
      ta.first_stage = first; ta.last_stage = last;
      ta.rin = rin; ta.rout = rout; ta.cin = cin; ta.cout = cout;
      ta.r_len = r_len; ta.c_len = c_len;
      (void)items;  // always one CTA per SM: narrow levels are dealt out one warp per SM first (msm_tail.cuh)
      // Complete projective formulas, two product rounds per addition: G1 on 8-lane teams ("tail_proj" >= 1), G2 on
      // whole warps — 8 roles x the 3 Karatsuba parts of an Fp2 product ("tail_proj" = 2; msm_tail_proj.cuh).
      if constexpr (std::is_same<F, Fp>::value) {
        if (opt_tail_proj) {
          B3 b3{3 * C::G1_B, 0};
          return launch_cooperative((const void*)msm_tail_proj<Fp>, 1u << 20, TAIL_THREADS, &ta, TAIL_SMEM_BYTES, &b3);
        }
      } else {
        if (opt_tail_proj >= 2) {
          B3 b3{3 * C::G2_B_RE, 3 * C::G2_B_IM};
          return launch_cooperative((const void*)msm_tail_proj<Fp2>, 1u << 20, TAIL_THREADS, &ta, TAIL_SMEM_BYTES, &b3);
        }
      }
      return launch_cooperative((const void*)msm_tail<F>, 1u << 20, TAIL_THREADS, &ta, TAIL_SMEM_BYTES);
    };
    const int tail_last = finish_in_tail ? TAIL_FINISH : TAIL_COMBINE;
    bool in_tail = false;
    // a bucket holds at most `count` records (count * W with the window table, where all windows share one
    // bucket space), i.e. at most ceil(that / chunk) + 1 tasks; fold levels beyond the largest task count actually
    // present (max_nt, on the device) are skipped
    const uint64_t items_bp = groups * (R * Lr + Cn * Lc);
    if (use_tail) {
      if (items_bp <= tail_cap) {
        rc = launch_tail(TAIL_FOLD, tail_last, 0, 0, items_bp);
        in_tail = true;
      } else {
        rc = launch_tail(TAIL_FOLD, TAIL_FOLD, 0, 0, 0);  // all fold levels in one launch
      }
      if (rc) return rc;
    } else {
      const uint64_t nt_bound = std::min<uint64_t>(max_tasks, count * (uint64_t)(precomp ? sh.W : 1) / (uint64_t)sh.chunk + 2);
      int fold_levels = 0;
      while ((1ull << fold_levels) < nt_bound && fold_levels < 31) fold_levels++;
      unsigned grid = (unsigned)std::min<uint64_t>((max_tasks + 127) / 128, (uint64_t)sms * 8);
      unsigned grid_team = (unsigned)std::min<uint64_t>((WIDE_LANES * max_tasks + 127) / 128, (uint64_t)sms * 8);
      for (int l = 0; l < fold_levels; l++) {
        if (opt_team)
          msm_fold_level_team<F><<<grid_team, 128, 0, stream>>>(l, task_first.as<uint32_t>() + nbuckets, d_hist + MSM_MAX_CHUNK + 1,
                                                                slot_seg.as<MsmSlotSeg>(), partial.as<XYZZ<F>>());
        else
          msm_fold_level<F><<<grid, 128, 0, stream>>>(l, task_first.as<uint32_t>() + nbuckets, d_hist + MSM_MAX_CHUNK + 1,
                                                      slot_seg.as<MsmSlotSeg>(), partial.as<XYZZ<F>>());
        LAUNCHED();
      }
    }
    // one level of the two trees as its own launch; false once the level fits the tail kernel
    auto wide_trees = [&](int stage, uint64_t sums0, uint64_t& len0, uint64_t sums1, uint64_t& len1) -> int {
      while (len0 > 1 || len1 > 1) {
        PairJob<F> j0{rin, rout, len0 > 1 ? sums0 * (len0 / 2) : 0};
        PairJob<F> j1{cin, cout, len1 > 1 ? sums1 * (len1 / 2) : 0};
        if (use_tail && j0.n_out + j1.n_out <= tail_cap) {
          in_tail = true;
          return launch_tail(stage, tail_last, len0, len1, j0.n_out + j1.n_out);
        }
        const uint64_t mx = std::max(j0.n_out, j1.n_out);
        if (opt_team && j0.n_out + j1.n_out <= TEAM_MAX_ITEMS) {  // latency-bound level: 4 lanes per addition
          dim3 grid((unsigned)((WIDE_LANES * mx + 127) / 128), 2);
          msm_pair_sum_team<F><<<grid, 128, 0, stream>>>(j0, j1);
        } else {
          dim3 grid((unsigned)std::min<uint64_t>((mx + 127) / 128, (uint64_t)148 * 64), 2);
          msm_pair_sum<F><<<grid, 128, 0, stream>>>(j0, j1);
        }
        LAUNCHED();
        if (len0 > 1) { std::swap(rin, rout); len0 /= 2; }
        if (len1 > 1) { std::swap(cin, cout); len1 /= 2; }
      }
      return KZG_GPU_OK;
    };
    if (!in_tail) {
      BucketJob<F> jr{rin, R, Cn, Cn, 1, groups * R * Lr};
      BucketJob<F> jc{cin, Cn, R, 1, Cn, groups * Cn * Lc};
      const uint64_t mx = std::max(jr.total, jc.total);
      if (opt_team && jr.total + jc.total <= TEAM_MAX_ITEMS) {
        dim3 grid((unsigned)((WIDE_LANES * mx + 127) / 128), 2);
        msm_bucket_pairs_team<F><<<grid, 128, 0, stream>>>((uint64_t)sh.nb, task_first.as<uint32_t>(), partial.as<XYZZ<F>>(), jr, jc);
      } else {
        dim3 grid((unsigned)std::min<uint64_t>((mx + 127) / 128, (uint64_t)sms * 64), 2);
        msm_bucket_pairs<F><<<grid, 128, 0, stream>>>((uint64_t)sh.nb, task_first.as<uint32_t>(), partial.as<XYZZ<F>>(), jr, jc);
      }
      LAUNCHED();
      uint64_t r_len = Lr, c_len = Lc;
      rc = wide_trees(TAIL_TREES1, groups * R, r_len, groups * Cn, c_len);
      if (rc) return rc;
    }
    if (!in_tail) {
      // rin = row sums [groups * R], cin = column sums [groups * Cn]; weighted chunks, then trees again
      WeightedJob<F> wr{rin, rout, R, groups * Pr, 0u};
      WeightedJob<F> wc{cin, cout, Cn, groups * Pc, 1u};
      if (use_tail && wr.total + wc.total <= tail_cap) {
        in_tail = true;
        rc = launch_tail(TAIL_WEIGHTED, tail_last, 0, 0, wr.total + wc.total);
        if (rc) return rc;
      } else {
        const uint64_t mw = std::max(wr.total, wc.total);
        if (opt_team && wr.total + wc.total <= TEAM_MAX_ITEMS) {
          dim3 gridw((unsigned)((WIDE_LANES * mw + 127) / 128), 2);
          msm_weighted_chunks_team<F><<<gridw, 128, 0, stream>>>(wr, wc);
        } else {
          dim3 gridw((unsigned)std::min<uint64_t>((mw + 127) / 128, (uint64_t)sms * 64), 2);
          msm_weighted_chunks<F><<<gridw, 128, 0, stream>>>(wr, wc);
        }
        LAUNCHED();
        std::swap(rin, rout);
        std::swap(cin, cout);
        uint64_t r_len = Pr, c_len = Pc;
        rc = wide_trees(TAIL_TREES2, groups, r_len, groups, c_len);
        if (rc) return rc;
      }
    }
    if (!in_tail) {
      const XYZZ<F>*t1 = rin, *t2 = cin;
      if (use_tail && groups <= tail_cap && finish_in_tail) {
        in_tail = true;
        rc = launch_tail(TAIL_COMBINE, tail_last, 0, 0, groups);
        if (rc) return rc;
      } else {
        if (opt_team && groups <= TEAM_MAX_ITEMS)
          msm_group_combine_team<F><<<(unsigned)((WIDE_LANES * groups + 127) / 128), 128, 0, stream>>>(groups, h, t1, t2, group_out.as<XYZZ<F>>());
        else
          msm_group_combine<F><<<(unsigned)((groups + 127) / 128), 128, 0, stream>>>(groups, h, t1, t2, group_out.as<XYZZ<F>>());
        LAUNCHED();
      }
    }
    phase_mark(3);
    if (in_tail && finish_in_tail) {
      // Horner over the windows, affine normalisation and bytes ran at the end of the tail kernel
    } else if (opt_team && sh.groups_per_msm > 1 && batch <= TEAM_MAX_ITEMS)  // Horner over the windows: a long dependent chain
      msm_finish_team<F><<<(unsigned)((TeamWidth<F>::LANES * batch + 127) / 128), 128, 0, stream>>>(
          batch, sh.groups_per_msm, sh.c, group_out.as<XYZZ<F>>(), out_bytes.as<uint8_t>(), out_inf.as<int>(), M,
          out_xyzz.as<XYZZ<F>>(), opt_finish_affine);
    else
    {
      const int lane_stride = batch <= 65536 ? 32 : 1;
      msm_finish<F><<<(unsigned)((batch * lane_stride + 127) / 128), 128, 0, stream>>>(
          batch, sh.groups_per_msm, sh.c, group_out.as<XYZZ<F>>(), out_bytes.as<uint8_t>(), out_inf.as<int>(), M,
          out_xyzz.as<XYZZ<F>>(), opt_finish_affine, lane_stride);
    }
    if (!(in_tail && finish_in_tail)) LAUNCHED();
    phase_mark(4);
    last_batch = batch;
    last_records = records; last_nbuckets = nbuckets; last_c = sh.c; last_chunk = sh.chunk;
    return KZG_GPU_OK;
  }

  // shape of a direct-lookup pass: ~2^18 threads when there is enough work, between 4 and 8 W additions per thread
  DirectShape direct_shape(uint64_t count, uint64_t batch, uint64_t base_offset) const {
    DirectShape sh{};
    sh.c = direct_c; sh.W = direct_W; sh.H = 1ull << (direct_c - 1); sh.n_table = srs_n;
    sh.n = count; sh.batch = batch; sh.base_offset = base_offset;
    const uint64_t adds = count * batch * (uint64_t)direct_W;
    const uint64_t per_thread = std::min<uint64_t>(8ull * direct_W, std::max<uint64_t>(4, adds >> 18));
    uint64_t spt = 1;
    if (per_thread >= (uint64_t)direct_W) { sh.wpt = direct_W; spt = per_thread / direct_W; }
    else sh.wpt = (int)per_thread;
    sh.groups = (direct_W + sh.wpt - 1) / sh.wpt;
    sh.spt = (int)spt;
    uint64_t Q = (count + spt - 1) / spt;
    if (sh.groups > 1) Q = (Q + 31) / 32 * 32;  // a warp stays inside one window group (direct_accumulate)
    sh.P = Q * (uint64_t)sh.groups;
    sh.P2 = 1;
    while (sh.P2 < sh.P) sh.P2 <<= 1;
    return sh;
  }
  // One direct-lookup pass for `batch` G1 MSMs of `count` scalars (device-resident bytes): accumulate, pairwise trees,
  // normalisation.  Same outputs as enqueue_msm.
  int enqueue_msm_direct(const uint8_t* d_scalars, uint64_t count, uint64_t batch, uint64_t base_offset) {
    using F = Fp;
    const DirectShape sh = direct_shape(count, batch, base_offset);
    const uint64_t slots = batch * sh.P2;
    CK(partial.ensure(slots * sizeof(XYZZ<F>)));
    CK(red0.ensure(std::max<uint64_t>(1, slots / 2) * sizeof(XYZZ<F>)));
    CK(red2.ensure(batch * sizeof(XYZZ<F>)));
    CK(red3.ensure(batch * sizeof(XYZZ<F>)));
    CK(group_out.ensure(batch * sizeof(XYZZ<F>)));
    CK(out_bytes.ensure(batch * (size_t)(4 * M)));
    CK(out_inf.ensure(batch * sizeof(int)));
    CK(out_xyzz.ensure(batch * sizeof(XYZZ<F>)));
    const bool use_tail = opt_tail && opt_team;
    if (use_tail) {
      int rct = prepare_tail<F>();
      if (rct) return rct;
    }
    for (int sl = 0; sl < chunk_slices; sl++) CK(cudaStreamWaitEvent(stream, chunk_ev[sl], 0));  // scalars still arriving in chunks
    chunk_slices = 0;
    phase_mark(0);
    phase_mark(1);
    direct_accumulate<C, F><<<(unsigned)std::min<uint64_t>((slots + 127) / 128, (uint64_t)1 << 30), 128, 0, stream>>>(
        sh, d_scalars, g1_direct.as<Affine<F>>(), partial.as<XYZZ<F>>());
    LAUNCHED();
    phase_mark(2);
    XYZZ<F>*in = partial.as<XYZZ<F>>(), *out = red0.as<XYZZ<F>>();
    uint64_t len = sh.P2;
    const bool finish_in_tail = use_tail && batch <= tail_threads_total() / 32;
    const uint64_t tail_cap = finish_in_tail ? 4 * tail_threads_total() : 0;
    while (len > 1 && batch * (len / 2) > tail_cap) {  // wide levels: launches of their own
      PairJob<F> j0{in, out, batch * (len / 2)}, j1{in, out, 0};
      if (opt_team && j0.n_out <= TEAM_MAX_ITEMS) {
        dim3 grid((unsigned)((WIDE_LANES * j0.n_out + 127) / 128), 1);
        msm_pair_sum_team<F><<<grid, 128, 0, stream>>>(j0, j1);
      } else {
        dim3 grid((unsigned)std::min<uint64_t>((j0.n_out + 127) / 128, (uint64_t)148 * 64), 1);
        msm_pair_sum<F><<<grid, 128, 0, stream>>>(j0, j1);
      }
      LAUNCHED();
      std::swap(in, out);
      len /= 2;
    }
    phase_mark(3);
    if (finish_in_tail) {
      // the rest in the cooperative tail kernel, entered at its second tree stage: "row sums" = our partial sums (len per
      // MSM), "column sums" = one point at infinity per MSM, no doublings in between (h = 0), then normalisation and bytes
      CK(cudaMemsetAsync(red2.p, 0, batch * sizeof(XYZZ<F>), stream));
      TailArgs<F> ta{};
      ta.first_stage = TAIL_TREES2; ta.last_stage = TAIL_FINISH;
      ta.barrier = barrier_buf.as<GridBarrier>();
      ta.rin = in; ta.rout = out; ta.cin = red2.as<XYZZ<F>>(); ta.cout = red3.as<XYZZ<F>>();
      ta.groups = batch; ta.R = 1; ta.Cn = 1; ta.r_len = len; ta.c_len = 1; ta.h = 0;
      ta.group_out = group_out.as<XYZZ<F>>();
      ta.batch = batch; ta.groups_per_msm = 1; ta.c = sh.c; ta.modbytes = M; ta.want_affine = opt_finish_affine;
      ta.out_bytes = out_bytes.as<uint8_t>(); ta.is_inf = out_inf.as<int>(); ta.out_xyzz = out_xyzz.as<XYZZ<F>>();
      int rc;
      if (opt_tail_proj) {
        B3 b3{3 * C::G1_B, 0};
        rc = launch_cooperative((const void*)msm_tail_proj<Fp>, 1u << 20, TAIL_THREADS, &ta, TAIL_SMEM_BYTES, &b3);
      } else {
        rc = launch_cooperative((const void*)msm_tail<Fp>, 1u << 20, TAIL_THREADS, &ta, TAIL_SMEM_BYTES);
      }
      if (rc) return rc;
    } else {
      // (large batches) `in` holds one sum per MSM
      if (opt_team && batch <= TEAM_MAX_ITEMS)
        msm_finish_team<F><<<(unsigned)((TeamWidth<F>::LANES * batch + 127) / 128), 128, 0, stream>>>(
            batch, 1, sh.c, in, out_bytes.as<uint8_t>(), out_inf.as<int>(), M, out_xyzz.as<XYZZ<F>>(), opt_finish_affine);
      else {
        const int lane_stride = batch <= 65536 ? 32 : 1;
        msm_finish<F><<<(unsigned)((batch * lane_stride + 127) / 128), 128, 0, stream>>>(
            batch, 1, sh.c, in, out_bytes.as<uint8_t>(), out_inf.as<int>(), M, out_xyzz.as<XYZZ<F>>(), opt_finish_affine, lane_stride);
      }
      LAUNCHED();
    }
    phase_mark(4);
    last_batch = batch;
    last_records = count * batch * (uint64_t)sh.W; last_nbuckets = 0; last_c = sh.c; last_chunk = sh.wpt;
    last_direct_partials = batch * sh.P;
    return KZG_GPU_OK;
  }

  // largest number of MSMs of `count` scalars that one pass may take
  uint64_t max_batch_per_pass(uint64_t count, int W) const {
    const uint64_t rec_budget = 1ull << 27;  // records per pass (digits + sorted = 1 GiB)
    uint64_t b = rec_budget / std::max<uint64_t>(1, count * (uint64_t)W);
    return std::max<uint64_t>(1, b);
  }

  int msm(int group, const uint8_t* scalars, bool on_device, uint64_t count, uint64_t batch, uint64_t base_offset,
          uint8_t* out, int* is_inf, bool do_fetch) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (srs_n == 0) return KZG_GPU_ERR_STATE;
    if (group == 2 && !have_g2) return KZG_GPU_ERR_STATE;
    if ((count && !scalars) || batch == 0) return KZG_GPU_ERR_ARG;
    if (base_offset + count > srs_n || base_offset + count < count) return KZG_GPU_ERR_ARG;
    last_group = group;
    if (group == 2) {
      int rc0 = ensure_g2_table(count);
      if (!rc0) rc0 = wait_g2_table();
      if (rc0) return rc0;
    }
    const int nco = group == 2 ? 4 : 2;
    if (count == 0) {  // empty sum = infinity (polyeval_G1 on the zero polynomial, src/trusted_setup.cpp:150-154)
      if (do_fetch) {
        if (out) memset(out, 0, batch * (size_t)(nco * M));
        if (is_inf) for (uint64_t b = 0; b < batch; b++) is_inf[b] = 1;
      }
      last_batch = 0;
      return KZG_GPU_OK;
    }
    const bool use_table = group == 1 ? g1_table.p != nullptr : g2_table.p != nullptr;
    const TableRung* rung = nullptr;  // G1: the smallest rung of the table ladder that holds every point of this MSM
    if (group == 1 && use_table && opt_table_ladder)
      for (const TableRung& rg : rungs)
        if (rg.n && base_offset + count <= rg.n) { rung = &rg; break; }
    const int Wguess = rung ? rung->W : use_table ? table_W : (C::SCALAR_BITS + 8) / 8;
    uint64_t per_pass = do_fetch ? max_batch_per_pass(count, Wguess) : batch;
    const bool direct = group == 1 && g1_direct.p != nullptr && opt_direct_table_gb > 0;
    if (direct && do_fetch)  // partial sums of a pass: at most 2 GiB
      per_pass = std::max<uint64_t>(1, std::min<uint64_t>(batch, (1ull << 31) / (direct_shape(count, batch, base_offset).P2 * sizeof(XYZZ<Fp>))));
    // Host scalars of a large batch: split into passes of >= 128 MSMs so that the copy of pass i + 1 (copy stream,
    // second staging buffer) runs under the kernels of pass i; one pass would leave the whole H2D exposed
    // (134 MB = 2.4 ms for 1024 blobs of 4096 coefficients in front of 35 ms of compute).
    const bool pipelined = !on_device && do_fetch && batch >= 256 && count * batch * (uint64_t)SB >= (32ull << 20);
    if (pipelined) per_pass = std::min<uint64_t>(per_pass, std::max<uint64_t>(128, (batch + 3) / 4));
    if (pipelined) {
      if (!copy_stream) CK(cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking));
      for (int i = 0; i < 2; i++) {
        if (!copied_ev[i]) CK(cudaEventCreateWithFlags(&copied_ev[i], cudaEventDisableTiming));
        if (!consumed_ev[i]) CK(cudaEventCreateWithFlags(&consumed_ev[i], cudaEventDisableTiming));
      }
    }
    DevBuf* stage_buf[2] = {&scal, &scal2};
    auto enqueue_copy = [&](uint64_t pass, uint64_t b0) -> int {  // pass -> staging buffer pass % 2, on the copy stream
      const uint64_t nb = std::min(per_pass, batch - b0);
      DevBuf* sb = stage_buf[pass & 1];
      if (pass >= 2) CK(cudaStreamWaitEvent(copy_stream, consumed_ev[pass & 1], 0));  // its previous user has finished
      CK(sb->ensure(per_pass * count * (size_t)SB));
      CK(cudaMemcpyAsync(sb->p, scalars + b0 * count * SB, nb * count * (size_t)SB, cudaMemcpyHostToDevice, copy_stream));
      CK(cudaEventRecord(copied_ev[pass & 1], copy_stream));
      return KZG_GPU_OK;
    };
    if (pipelined) {
      CK(scal.ensure(per_pass * count * (size_t)SB));   // both buffers at their final size before any copy is in flight
      CK(scal2.ensure(per_pass * count * (size_t)SB));
      int rc = enqueue_copy(0, 0);
      if (rc) return rc;
    }
    uint64_t pass = 0;
    for (uint64_t b0 = 0; b0 < batch; b0 += per_pass, pass++) {
      const uint64_t nb = std::min(per_pass, batch - b0);
      const uint8_t* d_scal;
      chunk_slices = 0;
      if (on_device) {
        d_scal = scalars + b0 * count * SB;
      } else if (pipelined) {
        CK(cudaStreamWaitEvent(stream, copied_ev[pass & 1], 0));
        d_scal = stage_buf[pass & 1]->as<uint8_t>();
      } else if (nb * count * (uint64_t)SB >= (8ull << 20) && nb * count >= 4 * 65536) {
        // one large host MSM: the scalars go over in 4 chunks on the copy stream and the digit kernel starts on
        // each chunk as it lands, so only the last quarter's digits are left after the copy (enqueue_msm, slices)
        CK(scal.ensure(nb * count * (size_t)SB));
        if (!copy_stream) CK(cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking));
        for (int sl = 0; sl < 4; sl++)
          if (!chunk_ev[sl]) CK(cudaEventCreateWithFlags(&chunk_ev[sl], cudaEventDisableTiming));
        CK(cudaEventRecord(chunk_ev[0], stream));              // the staging buffer's previous readers are done
        CK(cudaStreamWaitEvent(copy_stream, chunk_ev[0], 0));
        const uint64_t total = nb * count;
        for (int sl = 0; sl < 4; sl++) {
          const uint64_t t0 = total * sl / 4, t1 = total * (sl + 1) / 4;
          CK(cudaMemcpyAsync(scal.as<uint8_t>() + t0 * SB, scalars + (b0 * count + t0) * SB, (t1 - t0) * (size_t)SB,
                             cudaMemcpyHostToDevice, copy_stream));
          CK(cudaEventRecord(chunk_ev[sl], copy_stream));
        }
        chunk_slices = 4;
        d_scal = scal.as<uint8_t>();
      } else {
        CK(scal.ensure(nb * count * (size_t)SB));
        CK(cudaMemcpyAsync(scal.p, scalars + b0 * count * SB, nb * count * (size_t)SB, cudaMemcpyHostToDevice, stream));
        d_scal = scal.as<uint8_t>();
      }
      int rc;
      if (direct)
        rc = enqueue_msm_direct(d_scal, count, nb, base_offset);
      else if (group == 1)
        rc = rung ? enqueue_msm<Fp>(g1.as<Affine<Fp>>(), rung->buf.template as<Affine<Fp>>(), rung->c, d_scal, count, nb, base_offset, false, rung->n)
                  : enqueue_msm<Fp>(g1.as<Affine<Fp>>(), use_table ? g1_table.as<Affine<Fp>>() : nullptr, table_c, d_scal,
                                    count, nb, base_offset);
      else
        rc = enqueue_msm<Fp2>(g2.as<Affine<Fp2>>(), use_table ? g2_table.as<Affine<Fp2>>() : nullptr, table_c, d_scal,
                              count, nb, base_offset);
      if (rc) return rc;
      if (pipelined) {
        CK(cudaEventRecord(consumed_ev[pass & 1], stream));
        if (b0 + per_pass < batch) {  // the next pass's copy goes out while this pass computes
          rc = enqueue_copy(pass + 1, b0 + per_pass);
          if (rc) return rc;
        }
      }
      if (do_fetch) {
        rc = fetch_locked(out ? out + b0 * (size_t)(nco * M) : nullptr, is_inf ? is_inf + b0 : nullptr);
        if (rc) return rc;
      }
    }
    return KZG_GPU_OK;
  }

  int fetch_locked(uint8_t* out, int* is_inf) {
    const int nco = last_group == 2 ? 4 : 2;
    if (last_batch && out)
      CK(cudaMemcpyAsync(out, out_bytes.p, last_batch * (size_t)(nco * M), cudaMemcpyDeviceToHost, stream));
    if (last_batch && is_inf)
      CK(cudaMemcpyAsync(is_inf, out_inf.p, last_batch * sizeof(int), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    return KZG_GPU_OK;
  }
  int fetch(uint8_t* out, int* is_inf) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    return fetch_locked(out, is_inf);
  }

  // partial sums for the point-range split (G1 only): XYZZ of the last single MSM
  int fetch_partial(uint32_t* words, size_t* n_words) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (n_words) *n_words = sizeof(XYZZ<Fp>) / 4;
    if (!words) return KZG_GPU_OK;
    if (last_batch != 1 || last_group != 1) return KZG_GPU_ERR_STATE;
    CK(cudaMemcpyAsync(words, out_xyzz.p, sizeof(XYZZ<Fp>), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    return KZG_GPU_OK;
  }
  // device-resident variants for the NCCL path: the partial never visits the host
  int partial_device(void** d_ptr, size_t* bytes) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    if (!d_ptr || !bytes) return KZG_GPU_ERR_ARG;
    if (last_batch != 1 || last_group != 1 || !out_xyzz.p) return KZG_GPU_ERR_STATE;
    *d_ptr = out_xyzz.p;
    *bytes = sizeof(XYZZ<Fp>);
    return KZG_GPU_OK;
  }
  int sum_partials_device(const void* d_words, uint64_t n, uint8_t* out, int* is_inf) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (!d_words || n == 0 || n > 4096) return KZG_GPU_ERR_ARG;
    CK(out_bytes.ensure(4 * M));
    CK(out_inf.ensure(sizeof(int)));
    sum_points_team<Fp><<<1, 512, 0, stream>>>(n, static_cast<const XYZZ<Fp>*>(d_words), out_bytes.as<uint8_t>(), out_inf.as<int>(), M);
    LAUNCHED();
    last_batch = 1;
    last_group = 1;
    return fetch_locked(out, is_inf);
  }
  int sum_partials(const uint32_t* words, uint64_t n, uint8_t* out, int* is_inf) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (!words || n == 0 || n > 4096) return KZG_GPU_ERR_ARG;
    CK(partial.ensure(n * sizeof(XYZZ<Fp>)));
    CK(out_bytes.ensure(4 * M));
    CK(out_inf.ensure(sizeof(int)));
    CK(cudaMemcpyAsync(partial.p, words, n * sizeof(XYZZ<Fp>), cudaMemcpyHostToDevice, stream));
    sum_points_team<Fp><<<1, 512, 0, stream>>>(n, partial.as<XYZZ<Fp>>(), out_bytes.as<uint8_t>(), out_inf.as<int>(), M);
    LAUNCHED();
    last_batch = 1;
    last_group = 1;
    return fetch_locked(out, is_inf);
  }

  // ------------------------------------------------------------------ Fr polynomials
  // coefficients (host bytes) -> poly0 (Montgomery).  Returns with the copy + conversion enqueued.
  int upload_poly(const uint8_t* coeffs, uint64_t n) {
    CK(stage.ensure(n * (size_t)SB));
    CK(poly0.ensure(n * sizeof(Fr)));
    CK(poly1.ensure(n * sizeof(Fr)));
    CK(cudaMemcpyAsync(stage.p, coeffs, n * (size_t)SB, cudaMemcpyHostToDevice, stream));
    unsigned grid = (unsigned)std::min<uint64_t>((n + 127) / 128, 148 * 8);
    fr_from_bytes<C><<<grid, 128, 0, stream>>>(n, stage.as<uint8_t>(), poly0.as<Fr>());
    LAUNCHED();
    return KZG_GPU_OK;
  }
  // quotient of poly0 by the k consecutive roots -> qbytes (device, (n-k)*SB bytes)
  // quotients of `batch` polynomials (poly0, stride n) by k consecutive roots each -> qbytes (batch * (n-k) records).
  // d_roots: device array of the first roots (batched form) or nullptr (single polynomial, first_root).
  // h_roots: the same first roots on the host (lets a batch opened at few distinct positions share Z and 1 / rev(Z)).
  int enqueue_quotient(uint64_t n, int64_t first_root, uint64_t k, uint64_t batch = 1, const long long* d_roots = nullptr,
                       const int64_t* h_roots = nullptr) {
    const Fr* q;
    // measured on B200 at degree 4095 (tools/gpu_probe_quotient.py): the scan kernel costs 0.08 ms per root, the
    // power-series path 0.69 ms whatever k is -> series from 8 roots on
    // One polynomial: pick the cheapest of  k rounds on one CTA (~70 us + 1.4 ns per coefficient, per root),
    // k rounds over the whole GPU (~55 us + 0.02 ns per coefficient, per root: long polynomials, few roots) and the
    // power-series path (~0.3 ms + L^2 / 42 us for L = n - k quotient coefficients, whatever k is), from the timings
    // of tools/gpu_probe_quotient.py.  Batches keep one CTA per polynomial (the batch fills the machine).
    const uint64_t Lq = n - k;
    const bool single = batch == 1 && !d_roots;
    const double cost_cta = (double)k * (70.0 + (double)n * 1.4e-3);
    const double cost_grid = single && n >= 32768 ? (double)k * (55.0 + (double)n * 2e-5) : 1e300;
    const double cost_series = single && opt_long_division >= 3 && k >= 8 && k <= 65536 && Lq <= 65536
                                   ? 300.0 + (double)Lq * (double)Lq / 42e3 : 1e300;
    const bool series = cost_series < cost_cta && cost_series < cost_grid;
    const bool rounds_grid = !series && cost_grid < cost_cta && opt_long_division >= 3;
    // A batch with many roots: the systolic kernel runs one CTA per polynomial at ~184 M products/s (k L products each,
    // ~2 CTAs per SM), whatever the batch size.  When the polynomials are opened at few distinct positions — the usual
    // case: every blob at the same points — Z and the inverse series of rev(Z) are computed once per position over the
    // whole GPU and each polynomial only needs the last truncated product (L^2 / 2 products at the rate of the whole
    // GPU, ~268 products/ns): cheaper for every batch size, and it fills the machine for small batches.
    std::vector<long long> positions;
    bool batch_series = false;
    if (!single && d_roots && h_roots && opt_long_division >= 3 && k >= 48 && k <= 65536 && Lq <= 65536) {
      positions.assign(h_roots, h_roots + batch);
      std::sort(positions.begin(), positions.end());
      positions.erase(std::unique(positions.begin(), positions.end()), positions.end());
      const double cost_shared = (double)positions.size() * (310.0 + (double)Lq * (double)Lq / 42e3) +
                                 (double)batch * (double)Lq * (double)Lq / 268e3;
      const double cost_systolic = std::ceil((double)batch / 296.0) * ((double)k * (double)Lq / 184.0);
      batch_series = positions.size() <= 64 && cost_shared < cost_systolic;
    }
    // g = rev(Z)^-1 mod x^L for the k consecutive roots from `root` on (Newton doubling over truncated products)
    auto grid_for = [](uint64_t outs) { return (unsigned)std::min<uint64_t>((outs + 3) / 4, 148 * 32); };
    auto inverse_series = [&](int64_t root, Fr** g_out) -> int {
      Fr *pz = nullptr, *zz = nullptr;
      int rc = enqueue_interpolate(root, k, &pz, &zz, false);  // Z (k + 1 coefficients, monic)
      if (rc) return rc;
      const uint64_t L = n - k;
      for (DevBuf* bf : {&nt_g0, &nt_g1, &nt_u}) CK(bf->ensure(L * sizeof(Fr)));
      Fr *g = nt_g0.as<Fr>(), *g2 = nt_g1.as<Fr>(), *u = nt_u.as<Fr>();
      // g = 1 / rev(Z) mod x^1 = 1 (Z is monic); f = rev(Z) is read in place: f[t] = Z[k - t], only its first
      // min(k + 1, 2m) coefficients matter at precision 2m
      poly_mul_trunc<C><<<1, 128, 0, stream>>>(zz, 1, (long long)k, zz, 1, (long long)k, g, 1, 0, 0);  // 1 * 1
      LAUNCHED();
      for (uint64_t m = 1; m < L; m <<= 1) {
        const uint64_t m2 = std::min<uint64_t>(2 * m, L), nf = std::min<uint64_t>(k + 1, m2);
        poly_mul_trunc<C><<<grid_for(m2), 128, 0, stream>>>(zz, nf, (long long)k, g, m, -1, u, m2, 1, 0);   // u = 2 - f g
        LAUNCHED();
        poly_mul_trunc<C><<<grid_for(m2), 128, 0, stream>>>(g, m, -1, u, m2, -1, g2, m2, 0, 0);              // g' = g u
        LAUNCHED();
        std::swap(g, g2);
      }
      *g_out = g;
      return KZG_GPU_OK;
    };
    if (rounds_grid) {
      int rcb = ensure_barrier();
      if (rcb) return rcb;
      constexpr int QG = 256;
      CK(tile_sums.ensure(2 * 1024 * sizeof(Fr)));
      struct { uint64_t n; long long first_root; uint64_t k; Fr* b0; Fr* b1; Fr* agg; GridBarrier* bar; } qa{
          n, (long long)first_root, k, poly0.as<Fr>(), poly1.as<Fr>(), tile_sums.as<Fr>(), barrier_buf.as<GridBarrier>()};
      void* params[7] = {&qa.n, &qa.first_root, &qa.k, &qa.b0, &qa.b1, &qa.agg, &qa.bar};
      const void* kern = (const void*)poly_quotient_rounds_grid<C, QG>;
      int cap = max_cooperative_blocks(kern, QG);
      if (cap <= 0) { last_error = "cooperative launch: kernel does not fit the device"; return KZG_GPU_ERR_CUDA; }
      cap = std::min(std::min(cap, num_sms > 0 ? num_sms : cap), QG);
      CK(cudaLaunchCooperativeKernel(kern, dim3((unsigned)cap), dim3(QG), params, 0, stream));
      launches++;
      q = (k & 1) ? poly1.as<Fr>() : poly0.as<Fr>();
    } else if ((k < 48 || (single && opt_long_division >= 3)) && !series) {  // one CTA-wide suffix scan per root
      constexpr int QT = 512;
      poly_quotient_rounds<C, QT><<<(unsigned)batch, QT, 0, stream>>>(n, (long long)first_root, d_roots, k, poly0.as<Fr>(),
                                                                      poly1.as<Fr>());
      LAUNCHED();
      q = (k & 1) ? poly1.as<Fr>() : poly0.as<Fr>();
    } else if (series) {
      // one polynomial, many roots: rev(q) = rev(P) * rev(Z)^-1 mod x^L, all truncated products over the whole GPU
      Fr* g = nullptr;
      int rc = inverse_series(first_root, &g);
      if (rc) return rc;
      const uint64_t L = n - k;
      // rev(q) = rev(P)[0..L) * g mod x^L, written back to front: q[i] = rev(q)[L - 1 - i]
      poly_mul_trunc<C><<<grid_for(L), 128, 0, stream>>>(poly0.as<Fr>(), L, (long long)(n - 1), g, L, -1, poly1.as<Fr>(), L, 0, 1);
      LAUNCHED();
      q = poly1.as<Fr>();
    } else if (batch_series) {
      const uint64_t L = n - k;
      const unsigned gx = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>((L + 3) / 4, (148 * 32 + batch - 1) / batch));
      for (long long pos : positions) {
        Fr* g = nullptr;
        int rc = inverse_series(pos, &g);
        if (rc) return rc;
        for (uint64_t y0 = 0; y0 < batch; y0 += 65535) {  // gridDim.y limit
          const uint64_t ny = std::min<uint64_t>(65535, batch - y0);
          poly_mul_trunc<C><<<dim3(gx, (unsigned)ny), 128, 0, stream>>>(
              poly0.as<Fr>() + y0 * n, L, (long long)(n - 1), g, L, -1, poly1.as<Fr>() + y0 * n, L, 0, 1, n, n,
              positions.size() > 1 ? d_roots + y0 : nullptr, pos);
          LAUNCHED();
        }
      }
      q = poly1.as<Fr>();
    } else if (batch == 1 && !d_roots && k <= 4096 && opt_long_division) {
      // one polynomial, many roots: Z from the product tree (no P part), then CTA-wide long division
      Fr *pz = nullptr, *zz = nullptr;
      int rc = enqueue_interpolate(first_root, k, &pz, &zz, false);
      if (rc) return rc;
      const uint32_t kk = (uint32_t)k;
      const size_t smem = (size_t)(kk + 2) * sizeof(Fr);
      if (opt_long_division >= 2 && kk > 256) {  // 8-CTA cluster, one window element per thread (a cluster step costs
                                                 // ~1.1 us whatever k is; below 256 roots one CTA steps faster)
        const unsigned threads = std::max(32u, ((kk + DIV_CLUSTER - 1) / DIV_CLUSTER + 31) / 32 * 32);
        CK(cudaFuncSetAttribute(poly_long_division_cluster<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        poly_long_division_cluster<C><<<DIV_CLUSTER, threads, smem, stream>>>(n, kk, poly0.as<Fr>(), zz, poly1.as<Fr>());
      } else {
        const int ept = kk <= 1024 ? 1 : kk <= 2048 ? 2 : 4;
        const unsigned threads = ept == 1 ? ((kk + 31) / 32) * 32 : 1024;
        auto launch = [&](auto kern) -> int {
          CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
          kern<<<1, threads, smem, stream>>>(n, kk, poly0.as<Fr>(), zz, poly1.as<Fr>());
          return KZG_GPU_OK;
        };
        rc = ept == 1 ? launch(poly_long_division<C, 1>) : ept == 2 ? launch(poly_long_division<C, 2>) : launch(poly_long_division<C, 4>);
        if (rc) return rc;
      }
      CK(cudaGetLastError());
      LAUNCHED();
      q = poly1.as<Fr>();
    } else {       // many: systolic passes of up to ST roots, ping-ponging between the two buffers
      constexpr int ST = 512, CHK = 256;
      Fr* src = poly0.as<Fr>();
      Fr* dst = poly1.as<Fr>();
      uint64_t m = n, done = 0;
      while (done < k) {
        const int t = (int)std::min<uint64_t>(ST, k - done);
        poly_quotient_systolic<C, ST, CHK><<<(unsigned)batch, ST, 0, stream>>>(
            m, n, (long long)first_root + (long long)done, d_roots, (long long)done, t, src, dst);
        LAUNCHED();
        std::swap(src, dst);
        m -= t;
        done += t;
      }
      q = src;
    }
    const uint64_t nq = n - k;
    CK(qbytes.ensure(nq * batch * (size_t)SB));
    unsigned grid = (unsigned)std::min<uint64_t>((nq * batch + 127) / 128, 148 * 8);
    fr_to_bytes_strided<C><<<grid, 128, 0, stream>>>(nq, batch, n, q, qbytes.as<uint8_t>());
    LAUNCHED();
    return KZG_GPU_OK;
  }

  int poly_quotient(const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* q_out) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (!coeffs || !q_out || n == 0 || k >= n) return KZG_GPU_ERR_ARG;
    int rc = upload_poly(coeffs, n);
    if (rc) return rc;
    rc = enqueue_quotient(n, first_root, k);
    if (rc) return rc;
    CK(cudaMemcpyAsync(q_out, qbytes.p, (n - k) * (size_t)SB, cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    return KZG_GPU_OK;
  }

  int poly_eval(const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* y_out) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (!y_out || k == 0 || (n && !coeffs)) return KZG_GPU_ERR_ARG;
    if (n == 0) { memset(y_out, 0, k * (size_t)SB); return KZG_GPU_OK; }
    int rc = upload_poly(coeffs, n);
    if (rc) return rc;
    CK(poly1.ensure(k * sizeof(Fr)));
    CK(qbytes.ensure(k * (size_t)SB));
    constexpr int ET = 128;
    unsigned grid = (unsigned)std::min<uint64_t>(k, 148 * 16);
    poly_eval_range_kernel<C, ET><<<grid, ET, 0, stream>>>(n, (long long)first_root, k, poly0.as<Fr>(), poly1.as<Fr>());
    LAUNCHED();
    fr_to_bytes<C><<<(unsigned)std::min<uint64_t>((k + 127) / 128, 148 * 8), 128, 0, stream>>>(k, poly1.as<Fr>(), qbytes.as<uint8_t>());
    LAUNCHED();
    CK(cudaMemcpyAsync(y_out, qbytes.p, k * (size_t)SB, cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    return KZG_GPU_OK;
  }

  // create_proof: quotient stays on device and feeds the MSM directly
  int prove(const uint8_t* coeffs, uint64_t n, int64_t first_root, uint64_t k, uint8_t* out, int* is_inf) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (srs_n == 0) return KZG_GPU_ERR_STATE;
    if ((n && !coeffs) || k == 0) return KZG_GPU_ERR_ARG;
    if (k >= n) {  // zero quotient -> point at infinity
      if (out) memset(out, 0, 2 * M);
      if (is_inf) *is_inf = 1;
      return KZG_GPU_OK;
    }
    if (n - k > srs_n) return KZG_GPU_ERR_ARG;
    int rc = upload_poly(coeffs, n);
    if (rc) return rc;
    rc = enqueue_quotient(n, first_root, k);
    if (rc) return rc;
    // same stream: the MSM is ordered after the quotient; qbytes is not touched by the MSM path
    return msm(1, qbytes.as<uint8_t>(), true, n - k, 1, 0, out, is_inf, true);
  }

  // `batch` proofs in one call: quotients (one CTA per polynomial) feed the batched MSM without leaving the device
  int prove_batch(const uint8_t* coeffs, uint64_t n, const int64_t* first_roots, uint64_t k, uint64_t batch, uint8_t* out,
                  int* is_inf) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (srs_n == 0) return KZG_GPU_ERR_STATE;
    if (!coeffs || !first_roots || k == 0 || batch == 0 || n == 0) return KZG_GPU_ERR_ARG;
    if (k >= n) {
      if (out) memset(out, 0, batch * (size_t)(2 * M));
      if (is_inf) for (uint64_t b = 0; b < batch; b++) is_inf[b] = 1;
      return KZG_GPU_OK;
    }
    if (n - k > srs_n) return KZG_GPU_ERR_ARG;
    const uint64_t per_pass = std::max<uint64_t>(1, (1ull << 26) / n);  // coefficients per pass
    for (uint64_t b0 = 0; b0 < batch; b0 += per_pass) {
      const uint64_t nb = std::min(per_pass, batch - b0);
      int rc = upload_poly(coeffs + b0 * n * (size_t)SB, n * nb);
      if (rc) return rc;
      CK(roots.ensure(nb * sizeof(long long)));
      CK(cudaMemcpyAsync(roots.p, first_roots + b0, nb * sizeof(long long), cudaMemcpyHostToDevice, stream));
      rc = enqueue_quotient(n, 0, k, nb, roots.as<long long>(), first_roots + b0);
      if (rc) return rc;
      rc = msm(1, qbytes.as<uint8_t>(), true, n - k, nb, 0, out ? out + b0 * (size_t)(2 * M) : nullptr,
               is_inf ? is_inf + b0 : nullptr, true);
      if (rc) return rc;
    }
    return KZG_GPU_OK;
  }

  // interpolation through the consecutive nodes first_x .. first_x + k - 1 of the k values in `stage` (device, SB-byte
  // little-endian each); leaves the interpolant in *p_out (k coefficients) and the vanishing polynomial in *z_out (k + 1)
  int enqueue_interpolate(int64_t first_x, uint64_t k, Fr** p_out, Fr** z_out, bool with_p = true) {
    uint64_t kp = 1;
    while (kp < k) kp <<= 1;
    if (with_p && inv_fact_n < k) {  // 1/t! table (only the interpolant needs it), cached
      const uint64_t want = std::max<uint64_t>(kp, 4096);
      CK(inv_fact.ensure(want * sizeof(Fr)));
      fr_inv_factorials<C><<<1, INV_FACT_THREADS, 0, stream>>>((uint32_t)want, inv_fact.as<Fr>());
      LAUNCHED();
      inv_fact_n = want;
    }
    CK(it_u.ensure(k * sizeof(Fr)));
    CK(it_pa.ensure(kp * sizeof(Fr)));
    CK(it_pb.ensure(kp * sizeof(Fr)));
    CK(it_za.ensure(2 * kp * sizeof(Fr)));
    CK(it_zb.ensure(2 * kp * sizeof(Fr)));
    CK(qbytes.ensure((kp + 1) * (size_t)SB));
    if (with_p) {
      interp_scale_values<C><<<(unsigned)std::min<uint64_t>((k + 127) / 128, 148 * 8), 128, 0, stream>>>(
          k, stage.as<uint8_t>(), inv_fact.as<Fr>(), it_u.as<Fr>());
      LAUNCHED();
    }
    Fr *pin = it_pa.as<Fr>(), *pout = it_pb.as<Fr>(), *zin = it_za.as<Fr>(), *zout = it_zb.as<Fr>();
    interp_leaves<C><<<(unsigned)std::min<uint64_t>((kp + 3) / 4, 148 * 16), 128, 0, stream>>>(
        k, kp, (long long)first_x, it_u.as<Fr>(), inv_fact.as<Fr>(), pin, zin, with_p ? 1 : 0);
    LAUNCHED();
    for (uint64_t m = 1; m < kp; m <<= 1) {
      const uint64_t npairs = kp / (2 * m), warps = npairs * ((with_p ? 4 : 2) * m + 1);
      interp_tree_level<C><<<(unsigned)std::min<uint64_t>((warps + 3) / 4, 148 * 32), 128, 0, stream>>>(m, npairs, pin, zin, pout, zout,
                                                                                                        with_p ? 1 : 0);
      LAUNCHED();
      std::swap(pin, pout);
      std::swap(zin, zout);
    }
    *p_out = pin;
    *z_out = zin;
    return KZG_GPU_OK;
  }

  // polyfit / linear_roots_and_polyfit for a blob's consecutive nodes
  int interpolate(const uint8_t* ys, int64_t first_x, uint64_t k, uint8_t* coeffs_out, uint8_t* vanish_out) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (!ys || k == 0 || k > (1ull << 22) || (!coeffs_out && !vanish_out)) return KZG_GPU_ERR_ARG;
    CK(stage.ensure(k * (size_t)SB));
    CK(cudaMemcpyAsync(stage.p, ys, k * (size_t)SB, cudaMemcpyHostToDevice, stream));
    Fr *pin = nullptr, *zin = nullptr;
    int rc = enqueue_interpolate(first_x, k, &pin, &zin);
    if (rc) return rc;
    if (coeffs_out) {
      fr_to_bytes<C><<<(unsigned)std::min<uint64_t>((k + 127) / 128, 148 * 8), 128, 0, stream>>>(k, pin, qbytes.as<uint8_t>());
      LAUNCHED();
      CK(cudaMemcpyAsync(coeffs_out, qbytes.p, k * (size_t)SB, cudaMemcpyDeviceToHost, stream));
      CK(cudaStreamSynchronize(stream));
    }
    if (vanish_out) {
      fr_to_bytes<C><<<(unsigned)std::min<uint64_t>((k + 128) / 128, 148 * 8), 128, 0, stream>>>(k + 1, zin, qbytes.as<uint8_t>());
      LAUNCHED();
      CK(cudaMemcpyAsync(vanish_out, qbytes.p, (k + 1) * (size_t)SB, cudaMemcpyDeviceToHost, stream));
      CK(cudaStreamSynchronize(stream));
    }
    return KZG_GPU_OK;
  }

  // blob::from_bytes -> poly::from_blob -> create_commit chained on the device: raw bytes in, commitment out
  // (src/blob.cpp:19-48, src/poly.cpp:4-6, src/trusted_setup.cpp:137-147); only the document crosses the bus.
  int commit_bytes(const uint8_t* bytes, uint64_t byte_length, int chunk_size, int64_t chunk_offset, uint8_t* coeffs_out,
                   uint8_t* out, int* is_inf) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (srs_n == 0) return KZG_GPU_ERR_STATE;
    if (chunk_size < 1 || chunk_size > SB - 1 || byte_length % (uint64_t)chunk_size != 0) return KZG_GPU_ERR_ARG;
    const uint64_t k = byte_length / (uint64_t)chunk_size;
    if (k == 0) {
      if (out) memset(out, 0, 2 * M);
      if (is_inf) *is_inf = 1;
      return KZG_GPU_OK;
    }
    if (!bytes || k >= srs_n || k > (1ull << 22)) return KZG_GPU_ERR_ARG;  // create_commit: deg + 1 < _G1.size() (src/trusted_setup.cpp:138)
    CK(raw.ensure(byte_length));
    CK(stage.ensure(k * (size_t)SB));
    CK(cudaMemcpyAsync(raw.p, bytes, byte_length, cudaMemcpyHostToDevice, stream));
    blob_encode_chunks<<<(unsigned)std::min<uint64_t>((k * SB + 255) / 256, 148 * 16), 256, 0, stream>>>(
        k, chunk_size, SB, raw.as<uint8_t>(), stage.as<uint8_t>());
    LAUNCHED();
    Fr *pin = nullptr, *zin = nullptr;
    int rc = enqueue_interpolate(chunk_offset, k, &pin, &zin);
    if (rc) return rc;
    fr_to_bytes<C><<<(unsigned)std::min<uint64_t>((k + 127) / 128, 148 * 8), 128, 0, stream>>>(k, pin, qbytes.as<uint8_t>());
    LAUNCHED();
    if (coeffs_out) CK(cudaMemcpyAsync(coeffs_out, qbytes.p, k * (size_t)SB, cudaMemcpyDeviceToHost, stream));
    return msm(1, qbytes.as<uint8_t>(), true, k, 1, 0, out, is_inf, true);
  }

  // ------------------------------------------------------------------ trusted setup
  int ensure_fixed_base_tables() {
    if (fb_ready) return KZG_GPU_OK;
    std::lock_guard<std::mutex> guard(fb_mutex());
    FbShared& sh = fb_cache()[device];
    if (!sh.t2) {
      constexpr int W = fb_windows<C>();
      Affine<Fp>* t1 = nullptr;
      Affine<Fp2>* t2 = nullptr;
      CK(cudaMalloc(&t1, (size_t)W * FB_HALF * sizeof(Affine<Fp>)));
      cudaError_t e = cudaMalloc(&t2, (size_t)W * FB_HALF * sizeof(Affine<Fp2>));
      if (e != cudaSuccess) { cudaFree(t1); CK(e); }
      Affine<Fp> gen1;
      Affine<Fp2> gen2;
      for (int i = 0; i < Fp::L; i++) {
        gen1.x.v[i] = C::g1x(i); gen1.y.v[i] = C::g1y(i);
        gen2.x.a.v[i] = C::g2xa(i); gen2.x.b.v[i] = C::g2xb(i);
        gen2.y.a.v[i] = C::g2ya(i); gen2.y.b.v[i] = C::g2yb(i);
      }
      fixed_base_rows<C, Fp><<<(TeamWidth<Fp>::LANES * W + 127) / 128, 128, 0, stream>>>(gen1, t1);
      LAUNCHED();
      fixed_base_rows<C, Fp2><<<(TeamWidth<Fp2>::LANES * W + 127) / 128, 128, 0, stream>>>(gen2, t2);
      LAUNCHED();
      for (int level = 0; level + 1 < FB_C; level++) {
        const unsigned grid = (unsigned)(((uint64_t)W << level) + 127) / 128;
        fixed_base_fill<C, Fp><<<grid, 128, 0, stream>>>(level, t1);
        LAUNCHED();
        fixed_base_fill<C, Fp2><<<grid, 128, 0, stream>>>(level, t2);
        LAUNCHED();
      }
      e = cudaStreamSynchronize(stream);
      if (e == cudaSuccess) e = cudaGetLastError();
      if (e != cudaSuccess) { cudaFree(t1); cudaFree(t2); CK(e); }
      sh.t1 = t1;
      sh.t2 = t2;
    }
    fb1 = sh.t1;
    fb2 = sh.t2;
    fb_ready = true;
    return KZG_GPU_OK;
  }

  // Zero the buffers that held the secret and its powers (the reference frees s_powers when its constructor returns,
  // src/trusted_setup.cpp:21-74; here they would otherwise stay in HBM for the lifetime of the context).
  int scrub_secret(uint64_t n) {
    if (scal.p) cudaMemsetAsync(scal.p, 0, std::min<size_t>(scal.cap, 64), stream);
    if (powers.p) cudaMemsetAsync(powers.p, 0, std::min<size_t>(powers.cap, n * sizeof(Fr)), stream);
    if (fb_sums.p) cudaMemsetAsync(fb_sums.p, 0, fb_sums.cap, stream);
    cudaError_t e = cudaStreamSynchronize(stream);
    powers.release();
    if (fb_sums.cap > ((size_t)64 << 20)) fb_sums.release();  // 192 / 384 B per index: not kept next to a large SRS
    return e == cudaSuccess ? KZG_GPU_OK : cuda_fail(e, "scrub_secret");
  }
  int setup_powers(const uint8_t* s_le, uint64_t first_index, uint64_t n, uint8_t* g1_out, uint8_t* g2_out) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (!s_le || n == 0 || first_index + n < n) return KZG_GPU_ERR_ARG;
    const int rc_run = setup_powers_locked(s_le, first_index, n, g1_out, g2_out);
    const int rc_scrub = scrub_secret(n);  // on every exit
    if (rc_run) return rc_run;
    if (rc_scrub) return rc_scrub;
    srs_n = n;
    return build_table();
  }
  int setup_powers_locked(const uint8_t* s_le, uint64_t first_index, uint64_t n, uint8_t* g1_out, uint8_t* g2_out) {
    int rc = ensure_fixed_base_tables();
    if (rc) return rc;
    CK(stage.ensure(std::max<size_t>(n * (size_t)(4 * M), 64)));
    CK(powers.ensure(n * sizeof(Fr)));
    CK(g1.ensure(n * sizeof(Affine<Fp>)));
    CK(scal.ensure(64));
    CK(cudaMemcpyAsync(scal.p, s_le, SB, cudaMemcpyHostToDevice, stream));
    const int per_thread = 32;
    const uint64_t nthreads = (n + per_thread - 1) / per_thread;
    setup_powers_kernel<C><<<(unsigned)((nthreads + 63) / 64), 64, 0, stream>>>(n, first_index, scal.as<uint8_t>(), per_thread, powers.as<uint32_t>());
    LAUNCHED();
    const unsigned grid = (unsigned)((n + 127) / 128);
    const unsigned grid_aff = (unsigned)(((n + FB_BATCH - 1) / FB_BATCH + 127) / 128);
    const bool want_g2 = g2_out || opt_setup_g2;  // "setup_g2" = 0: G1-only SRS (large synthetic MSM benches)
    CK(fb_sums.ensure(n * (want_g2 ? sizeof(XYZZ<Fp2>) : sizeof(XYZZ<Fp>))));
    fixed_base_mul<C, Fp><<<grid, 128, 0, stream>>>(n, powers.as<uint32_t>(), fb1, fb_sums.as<XYZZ<Fp>>());
    LAUNCHED();
    fixed_base_to_affine<Fp><<<grid_aff, 128, 0, stream>>>(n, fb_sums.as<XYZZ<Fp>>(), g1.as<Affine<Fp>>(),
                                                           g1_out ? stage.as<uint8_t>() : nullptr, M);
    LAUNCHED();
    if (g1_out) CK(cudaMemcpyAsync(g1_out, stage.p, n * (size_t)(2 * M), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    if (want_g2) {
      CK(g2.ensure(n * sizeof(Affine<Fp2>)));
      fixed_base_mul<C, Fp2><<<grid, 128, 0, stream>>>(n, powers.as<uint32_t>(), fb2, fb_sums.as<XYZZ<Fp2>>());
      LAUNCHED();
      fixed_base_to_affine<Fp2><<<grid_aff, 128, 0, stream>>>(n, fb_sums.as<XYZZ<Fp2>>(), g2.as<Affine<Fp2>>(),
                                                              g2_out ? stage.as<uint8_t>() : nullptr, M);
      LAUNCHED();
      if (g2_out) CK(cudaMemcpyAsync(g2_out, stage.p, n * (size_t)(4 * M), cudaMemcpyDeviceToHost, stream));
      CK(cudaStreamSynchronize(stream));
    }
    have_g2 = want_g2;
    return KZG_GPU_OK;
  }

  // ------------------------------------------------------------------ diagnostics
  // CUDA-event durations of the last MSM pass ("profile" option): sort+tasks, accumulate, reduce, finish
  int phase_times(float* ms, int n) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (!ms || n < 4) return KZG_GPU_ERR_ARG;
    if (!opt_profile) return KZG_GPU_ERR_STATE;
    CK(cudaEventSynchronize(phase_ev[4]));
    for (int i = 0; i < 4; i++) CK(cudaEventElapsedTime(&ms[i], phase_ev[i], phase_ev[i + 1]));
    return KZG_GPU_OK;
  }
  // shape of the most recent MSM pass: records, buckets, tasks, most tasks in one bucket, window bits, task chunk
  uint64_t last_records = 0, last_nbuckets = 0;
  int last_c = 0, last_chunk = 0;
  uint64_t last_direct_partials = 0;
  int msm_stats(uint64_t* out, int n) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (!out || n < 6) return KZG_GPU_ERR_ARG;
    if (!last_nbuckets && last_records && last_c) {  // a direct-lookup pass: no buckets; "tasks" = partial sums, chunk = windows per thread
      out[0] = last_records; out[1] = 0; out[2] = last_direct_partials; out[3] = 0; out[4] = (uint64_t)last_c; out[5] = (uint64_t)last_chunk;
      return KZG_GPU_OK;
    }
    if (!last_nbuckets) return KZG_GPU_ERR_STATE;
    uint32_t ntasks = 0, max_nt = 0;
    CK(cudaMemcpyAsync(&ntasks, task_first.as<uint32_t>() + last_nbuckets, 4, cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(&max_nt, hist.as<uint32_t>() + MSM_MAX_CHUNK + 1, 4, cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    out[0] = last_records; out[1] = last_nbuckets; out[2] = ntasks; out[3] = max_nt; out[4] = (uint64_t)last_c; out[5] = (uint64_t)last_chunk;
    return KZG_GPU_OK;
  }
  int microbench(int kind, double* ops) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (!ops || kind < 0 || kind > 13) return KZG_GPU_ERR_ARG;
    CK(scal.ensure(64));
    if (kind >= 4) return microbench_latency(kind, ops);
    const int blocks = 148 * 8, threads = 256;
    int iters = kind <= 1 ? 4096 : 256;
    double per_thread_iter = kind == 0 ? 32.0 : kind == 1 ? 16.0 : 2.0;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
      CK(cudaEventRecord(e0, stream));
      switch (kind) {
        case 0: bench_wide_chain<<<blocks, threads, 0, stream>>>(iters, 12345u + rep, scal.as<uint32_t>()); break;
        case 1: bench_imad32<<<blocks, threads, 0, stream>>>(iters, 12345u + rep, scal.as<uint32_t>()); break;
        case 2: bench_modmul<Fp><<<blocks, threads, 0, stream>>>(iters, 12345u + rep, scal.as<uint32_t>()); break;
        default: bench_modmul<Fr><<<blocks, threads, 0, stream>>>(iters, 12345u + rep, scal.as<uint32_t>()); break;
      }
      LAUNCHED();
      CK(cudaEventRecord(e1, stream));
      CK(cudaEventSynchronize(e1));
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (rep > 0) best = std::min(best, ms);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *ops = (double)blocks * threads * iters * per_thread_iter / (best * 1e-3);
    return KZG_GPU_OK;
  }

  // kinds 4..7: latency of one grid barrier / team addition / single-thread addition / Montgomery product (G1 field), in
  // the msm_tail configuration (one CTA of 128 threads per SM).  *ops = steps per second of ONE dependent chain.
  int microbench_latency(int kind, double* ops) {
    int rc = prepare_tail<Fp>();
    if (rc) return rc;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int iters = kind == 4 ? 2000 : kind == 7 ? 4000 : kind >= 10 ? 400 : kind >= 8 ? 50 : 400;
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
      CK(cudaEventRecord(e0, stream));
      if (kind == 4) {
        int it = iters;
        GridBarrier* bar = barrier_buf.as<GridBarrier>();
        void* params[2] = {&it, &bar};
        const void* kern = (const void*)bench_grid_barrier;
        int cap = max_cooperative_blocks(kern, TAIL_THREADS);
        if (cap <= 0) return KZG_GPU_ERR_CUDA;
        cap = std::min(cap, num_sms > 0 ? num_sms : cap);
        CK(cudaLaunchCooperativeKernel(kern, dim3((unsigned)cap), dim3(TAIL_THREADS), params, 0, stream));
        launches++;
      } else {
        if (kind >= 10) {
          B3 b3{kind == 13 ? 3 * C::G1_B : 3 * C::G2_B_RE, kind == 13 ? 0 : 3 * C::G2_B_IM};
          bench_latency_chain_g2<Fp, Fp2><<<num_sms > 0 ? num_sms : 148, TAIL_THREADS, 0, stream>>>(kind, iters, 777u + rep, b3, scal.as<uint32_t>());
        } else {
          bench_latency_chain<Fp><<<num_sms > 0 ? num_sms : 148, TAIL_THREADS, 0, stream>>>(kind, iters, 777u + rep, scal.as<uint32_t>());
        }
        LAUNCHED();
      }
      CK(cudaEventRecord(e1, stream));
      CK(cudaEventSynchronize(e1));
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (rep > 0) best = std::min(best, ms);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *ops = (double)iters / (best * 1e-3);
    return KZG_GPU_OK;
  }

  int selftest(int* failures) override {
    std::lock_guard<std::recursive_mutex> lock(mu);
    CK(cudaSetDevice(device));
    if (!failures) return KZG_GPU_ERR_ARG;
    CK(out_inf.ensure(sizeof(int)));
    CK(cudaMemsetAsync(out_inf.p, 0, sizeof(int), stream));
    selftest_kernel<C><<<4, 32, 0, stream>>>(20261104u, out_inf.as<int>());
    LAUNCHED();
    CK(cudaMemcpyAsync(failures, out_inf.p, sizeof(int), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    return KZG_GPU_OK;
  }
};


// factories, one per translation unit
kzg_gpu_ctx* kzg_gpu_make_ctx_bn158(int device, int* rc);
kzg_gpu_ctx* kzg_gpu_make_ctx_bn254(int device, int* rc);
kzg_gpu_ctx* kzg_gpu_make_ctx_bls12381(int device, int* rc);

#define KZG_GPU_DEFINE_FACTORY(fn, CURVE)                         \
  kzg_gpu_ctx* fn(int device, int* rc) {                          \
    auto* p = new (std::nothrow) Ctx<kzgb200::CURVE>();           \
    if (!p) { *rc = KZG_GPU_ERR_MEMORY; return nullptr; }         \
    p->device = device;                                           \
    *rc = p->init();                                              \
    if (*rc) { delete p; return nullptr; }                        \
    return p;                                                     \
  }
