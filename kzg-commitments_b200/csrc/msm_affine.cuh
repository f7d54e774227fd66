// Batched-affine pre-reduction of the sorted bucket records (levels of pairwise AFFINE additions).
//
// msm_accumulate spends 10 modmuls per point (XYZZ mixed addition) and already keeps the integer
// multiply pipe ~90 % busy (profiles/r1e), so the only way to go faster is fewer multiplications.
// An affine addition costs 1 squaring + 2 products once 1/(x2 - x1) is known, and Montgomery's trick
// shares one inversion among any number of independent additions at 3 products each: 6 modmuls.
// One level replaces every bucket's records by the sums of adjacent pairs (an odd one out is copied):
//   level l output of bucket b:  ceil(cnt_b / 2^l) points at  off_l[b] + j   (off_l = scan of those sizes)
// Three launches per level:
//   aff_forward   each thread walks K consecutive output slots: loads the pair, forms the denominator
//                 d (x2 - x1; 2*y for P + P; 1 where no inversion is needed), stores the exclusive running
//                 product E_k = d_0 ... d_{k-1} and finally the thread's total product
//   aff_invert    inverts all totals: per warp, thread-local running products, shuffle scans, ONE
//                 inversion by lane 0 (binary GCD, no divergence), back-substitution
//   aff_backward  walks the same slots in reverse: 1/d_k = I_k * E_k, I_{k-1} = I_k * d_k, then the sum
// Exceptional inputs produce what the complete formulas of the reference (ecp.cpp:750-812) would:
// infinity operands ((0,0)), P + P (doubling, through the same shared inversion), P + (-P) = infinity.
// After the chosen number of levels the surviving points feed the XYZZ task pipeline of msm.cuh unchanged.
#pragma once
#include "ec.cuh"

namespace kzgb200 {

constexpr int AFF_K = 32;  // output slots per thread

enum AffKind : int { AFF_NONE = 0, AFF_COPY = 1, AFF_ADD = 2, AFF_DBL = 3, AFF_INF = 4, AFF_TAKE2 = 5 };

template <class F>
struct AffLevel {
  // input: either records into a base array (level 0) or the previous level's points
  const uint32_t* sorted;    // level 0: record = source index | sign << 31 ; nullptr afterwards
  const Affine<F>* in_pts;   // level 0: bases / window table ; afterwards: previous level's output
  const uint32_t* off_in;    // per-bucket start of the input segment (nbuckets + 1 entries)
  const uint32_t* off_out;   // per-bucket start of the output segment (nbuckets + 1 entries)
  uint64_t nbuckets;
  Affine<F>* out_pts;
  F* prefix;                 // E_k per output slot
  F* totals;                 // per thread
};

template <class F>
KZG_HD Affine<F> aff_load(const AffLevel<F>& lv, uint32_t i) {
  if (!lv.sorted) return lv.in_pts[i];
  const uint32_t rec = lv.sorted[i];
  Affine<F> p = lv.in_pts[rec & 0x7fffffffu];
  if (rec >> 31) p.y = F::neg(p.y);
  return p;
}

// bucket containing output slot p: largest b with off_out[b] <= p (empty buckets share their successor's offset)
template <class F>
KZG_HD uint64_t aff_bucket_of(const AffLevel<F>& lv, uint32_t p) {
  uint64_t lo = 0, hi = lv.nbuckets;  // invariant: off_out[lo] <= p < off_out[hi]
  while (hi - lo > 1) {
    const uint64_t mid = (lo + hi) >> 1;
    if (lv.off_out[mid] <= p) lo = mid; else hi = mid;
  }
  return lo;
}

// classify the pair behind output slot p of bucket b and return its denominator
template <class F>
KZG_HD int aff_pair(const AffLevel<F>& lv, uint64_t b, uint32_t p, Affine<F>& p1, Affine<F>& p2, F& d) {
  const uint32_t j = p - lv.off_out[b];
  const uint32_t in0 = lv.off_in[b], cnt = lv.off_in[b + 1] - in0;
  p1 = aff_load<F>(lv, in0 + 2 * j);
  if (2 * j + 1 >= cnt) return AFF_COPY;
  p2 = aff_load<F>(lv, in0 + 2 * j + 1);
  if (p2.is_inf()) return AFF_COPY;
  if (p1.is_inf()) return AFF_TAKE2;
  if (p1.x == p2.x) {
    if (p1.y == p2.y && !p1.y.is_zero()) { d = F::dbl(p1.y); return AFF_DBL; }
    return AFF_INF;
  }
  d = F::sub(p2.x, p1.x);
  return AFF_ADD;
}

template <class F>
KZG_HD void aff_forward_thread(const AffLevel<F>& lv, uint64_t t) {
  const uint32_t total = lv.off_out[lv.nbuckets];
  const uint64_t p0 = t * AFF_K;
  if (p0 >= total) return;
  const uint32_t p1e = (uint32_t)((p0 + AFF_K < total) ? p0 + AFF_K : total);
  uint64_t b = aff_bucket_of<F>(lv, (uint32_t)p0);
  F run = F::one();
  for (uint32_t p = (uint32_t)p0; p < p1e; p++) {
    while (lv.off_out[b + 1] <= p) b++;
    Affine<F> a1, a2;
    F d;
    const int kind = aff_pair<F>(lv, b, p, a1, a2, d);
    lv.prefix[p] = run;
    if (kind == AFF_ADD || kind == AFF_DBL) run = F::mul(run, d);
  }
  lv.totals[t] = run;
}

template <class F>
KZG_HD void aff_backward_thread(const AffLevel<F>& lv, uint64_t t) {
  const uint32_t total = lv.off_out[lv.nbuckets];
  const uint64_t p0 = t * AFF_K;
  if (p0 >= total) return;
  const uint32_t p1e = (uint32_t)((p0 + AFF_K < total) ? p0 + AFF_K : total);
  uint64_t b = aff_bucket_of<F>(lv, p1e - 1);
  F inv_run = lv.totals[t];  // inverse of the thread's total product (after aff_invert)
  for (uint32_t p = p1e; p-- > (uint32_t)p0;) {
    while (lv.off_out[b] > p) b--;
    Affine<F> a1, a2, r;
    F d;
    const int kind = aff_pair<F>(lv, b, p, a1, a2, d);
    if (kind == AFF_COPY) r = a1;
    else if (kind == AFF_TAKE2) r = a2;
    else if (kind == AFF_INF) r = Affine<F>::inf();
    else {
      const F inv_d = F::mul(inv_run, lv.prefix[p]);
      inv_run = F::mul(inv_run, d);
      F lam;
      if (kind == AFF_ADD) lam = F::mul(F::sub(a2.y, a1.y), inv_d);
      else {
        const F xx = F::sqr(a1.x);
        lam = F::mul(F::add(F::dbl(xx), xx), inv_d);  // 3 x^2 / (2 y)
      }
      r.x = F::sub(F::sub(F::sqr(lam), a1.x), a2.x);   // for a doubling a2 == a1
      r.y = F::sub(F::mul(lam, F::sub(a1.x, r.x)), a1.y);
    }
    lv.out_pts[p] = r;
  }
}

#if defined(__CUDACC__)
template <class F>
__global__ void __launch_bounds__(128) aff_forward(AffLevel<F> lv, uint64_t nthreads) {
  const uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (t < nthreads) aff_forward_thread<F>(lv, t);
}
template <class F>
__global__ void __launch_bounds__(128) aff_backward(AffLevel<F> lv, uint64_t nthreads) {
  const uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (t < nthreads) aff_backward_thread<F>(lv, t);
}
#endif

}  // namespace kzgb200

#if defined(__CUDACC__)
namespace kzgb200 {

template <class P>
__device__ __forceinline__ Field<P> shfl_field(const Field<P>& v, int src) {
  Field<P> r;
#pragma unroll
  for (int i = 0; i < Field<P>::L; i++) r.v[i] = __shfl_sync(0xffffffffu, v.v[i], src);
  return r;
}

// totals[i] <- 1 / totals[i] for i < ceil(*total_slots / AFF_K).  One warp per 32*K2 totals: thread-local
// exclusive products, inclusive prefix/suffix scans over the lanes (shuffles), one inversion per warp.
template <class F, int K2>
__global__ void __launch_bounds__(128) aff_invert(const uint32_t* __restrict__ total_slots, F* __restrict__ totals) {
  const uint64_t n_tot = ((uint64_t)*total_slots + AFF_K - 1) / AFF_K;
  const int lane = threadIdx.x & 31;
  const uint64_t warp = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
  if (warp * 32 * K2 >= n_tot) return;  // whole warp out of range
  const uint64_t i0 = (warp * 32 + lane) * K2;
  F e[K2];
  F run = F::one();
#pragma unroll
  for (int k = 0; k < K2; k++) {
    e[k] = run;
    if (i0 + k < n_tot) run = F::mul(run, totals[i0 + k]);
  }
  F pfx = run, sfx = run;
  for (int d = 1; d < 32; d <<= 1) {
    const F up = shfl_field(pfx, lane >= d ? lane - d : lane);
    const F dn = shfl_field(sfx, lane + d < 32 ? lane + d : lane);
    if (lane >= d) pfx = F::mul(pfx, up);
    if (lane + d < 32) sfx = F::mul(sfx, dn);
  }
  F tinv = shfl_field(pfx, 31);
  if (lane == 0) tinv = F::inv(tinv);
  tinv = shfl_field(tinv, 0);
  const F before = shfl_field(pfx, lane > 0 ? lane - 1 : 0);
  const F after = shfl_field(sfx, lane < 31 ? lane + 1 : 31);
  F inv = tinv;
  if (lane > 0) inv = F::mul(inv, before);
  if (lane < 31) inv = F::mul(inv, after);
#pragma unroll
  for (int k = K2 - 1; k >= 0; k--) {
    if (i0 + k < n_tot) {
      const F v = totals[i0 + k];
      totals[i0 + k] = F::mul(inv, e[k]);
      inv = F::mul(inv, v);
    }
  }
}

}  // namespace kzgb200
#endif
