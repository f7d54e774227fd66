// MSM by DIRECT LOOKUP: the per-SRS window table taken one step further, for a SMALL SRS on a GPU with 180 GB.
//
// The window table of msm.cuh holds T[w][i] = 2^(c w) G_i, so a commitment is a bucket MSM over n W (point, digit)
// records: sort by digit, accumulate buckets, then reduce sum_d d * B_d — and for the SRS sizes the reference's callers
// use (4096-element blobs, src/blob.cpp; benchmark.cpp degree 4096) the reduction of 2^(c-1) buckets and the sort are
// what a single create_commit waits for (0.33 of 0.46 ms at 4097 terms, profiles/r2_tail.md).  Here the table holds
// every multiple a digit can select,
//     D[w][i][d-1] = d * 2^(c w) * G_i,      d = 1 .. 2^(c-1),
// so polyeval_G1 (src/trusted_setup.cpp:149-174) is a plain SUM of n W looked-up affine points (sign from the
// digit): no sort, no buckets, no weighting by the bucket index.  4097 points, c = 13: 20 x 4097 x 4096 entries of
// 96 bytes = 32 GB of HBM per SRS, built once in ~0.4 s — hence strictly opt-in ("direct_table_gb" option: the memory
// the table may take).
//
//   (entry d = 1)       D[w][i][0] = 2^(c w) G_i: written by msm_precompute_table with a row of H entries as its stride
//   direct_table_fill   level j doubles the filled prefix of every row: (2^j + r) B = 2^j B + r B, r = 1 .. 2^j
//                       (the same scheme as fixed_base_fill in setup.cuh, over W n rows instead of W)
//   direct_accumulate   one thread per (MSM, scalar, group of windows): signed digits of the scalar (as msm_digits), one
//                       gathered point and one mixed XYZZ addition per non-zero digit, one partial sum per thread
// The partial sums of each MSM are then added by the pairwise-tree steps that already exist (msm_pair_sum for wide
// levels, the cooperative tail kernel from TAIL_TREES2 on), which also normalise and write the bytes.
#pragma once
#include "msm.cuh"

namespace kzgb200 {

struct DirectShape {
  int c, W;              // window bits and windows per scalar of the table
  uint64_t H;            // entries per (window, point) row = 2^(c-1)
  uint64_t n_table;      // points per window of the table (the SRS size)
  uint64_t n;            // scalars per MSM
  uint64_t batch;        // MSMs
  uint64_t base_offset;  // first SRS point used
  int wpt;               // windows per thread
  int groups;            // window groups per scalar = ceil(W / wpt)
  int spt;               // scalars per thread (> 1 only when wpt = W)
  uint64_t P;            // partial sums per MSM = Q * groups, Q = ceil(n / spt) (rounded up to a multiple of 32 when groups > 1)
  uint64_t P2;           // slots per MSM (power of two >= P; the rest are written as infinity)
};

// The per-thread bodies are plain host/device functions (as in msm.cuh) so that csrc/host_check.cpp can run the same
// code on the CPU through the carry-flag emulation; the __global__ wrappers only add the grid-stride loop.
// entry (2^level + r) of a row, r = 1 .. 2^level (r = 2^level is the doubling: the mixed addition's P + P branch)
template <class F>
KZG_HD void direct_fill_thread(int level, uint64_t H, Affine<F>* direct, uint64_t t) {
  const uint64_t per = 1ull << level;
  const uint64_t row = t >> level, r = (t & (per - 1)) + 1;
  Affine<F>* e = direct + row * H;
  XYZZ<F> acc = XYZZ<F>::from_affine(e[per - 1]);
  acc.madd(e[r - 1], false);
  e[per + r - 1] = acc.is_inf() ? Affine<F>::inf() : acc.to_affine();
}

// one partial sum per slot: slot p of MSM b is (scalars q spt .. q spt + spt - 1, window group g), p = g * (P / groups) + q;
// slots P .. P2-1 are infinity.  The digits are those of msm_digits_thread (signed, carry from the windows below).
template <class C, class F>
KZG_HD void direct_accumulate_thread(const DirectShape& sh, const uint8_t* scalars, const Affine<F>* direct, XYZZ<F>* partial,
                                     uint64_t t) {
  const uint64_t b = t / sh.P2, p = t - b * sh.P2;
  XYZZ<F> acc = XYZZ<F>::inf();
  if (p < sh.P) {
    // group-major: the threads of a warp share the window group, so they reach their additions in the same loop
    // iterations (scalar-major slots made a warp walk through all W windows one divergent subset at a time: 4x slower)
    // (Q = P / groups is a multiple of 32 when there are several groups, so a warp never straddles two of them)
    const uint64_t Q = sh.P / (uint64_t)sh.groups;
    const int g = (int)(p / Q);
    const uint64_t q = p - (uint64_t)g * Q;
    const int w0 = g * sh.wpt, w1 = (w0 + sh.wpt < sh.W) ? w0 + sh.wpt : sh.W;
    const uint32_t half = 1u << (sh.c - 1);
    const uint64_t i_end = (q + 1) * (uint64_t)sh.spt;
    const uint64_t i1 = i_end < sh.n ? i_end : sh.n;
    for (uint64_t i = q * (uint64_t)sh.spt; i < i1; i++) {
      uint32_t k[C::FrP::L];
      load_scalar<C>(scalars, b * sh.n + i, k);
      uint32_t carry = 0;
      for (int w = 0; w < w1; w++) {
        const uint32_t raw = scalar_bits<C::FrP::L>(k, w * sh.c, sh.c) + carry;
        int32_t d;
        if (raw > half) { d = (int32_t)raw - (int32_t)(1u << sh.c); carry = 1; }
        else { d = (int32_t)raw; carry = 0; }
        if (w >= w0 && d != 0) {
          const uint32_t mag = (uint32_t)(d < 0 ? -d : d);
          const Affine<F> pt = direct[((uint64_t)w * sh.n_table + sh.base_offset + i) * sh.H + (mag - 1)];
          acc.madd(pt, d < 0);
        }
      }
    }
  }
  partial[t] = acc;
}

#if defined(__CUDACC__)
template <class F>
__global__ void __launch_bounds__(128)
direct_table_fill(int level, uint64_t rows, uint64_t H, Affine<F>* __restrict__ direct) {
  const uint64_t total = rows << level;
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < total; t += (uint64_t)gridDim.x * blockDim.x)
    direct_fill_thread<F>(level, H, direct, t);
}
template <class C, class F>
__global__ void __launch_bounds__(128)
direct_accumulate(DirectShape sh, const uint8_t* __restrict__ scalars, const Affine<F>* __restrict__ direct,
                  XYZZ<F>* __restrict__ partial) {
  const uint64_t total = sh.batch * sh.P2;
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < total; t += (uint64_t)gridDim.x * blockDim.x)
    direct_accumulate_thread<C, F>(sh, scalars, direct, partial, t);
}
#endif

}  // namespace kzgb200
