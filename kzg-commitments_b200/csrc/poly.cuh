// Fr polynomial kernels for create_proof (src/trusted_setup.cpp:214-228).
//
//   poly_quotient    q = P div Z,  Z = prod_{j<k} (x - (first_root + j))
//                    == the reference's  (P - I) / Z  (src/trusted_setup.cpp:225 -> NTL div,
//                    ntl/src/ZZ_pX.cpp:3953): I = P mod Z has degree < k, so the quotient is the same.
//                    Computed WITHOUT I, Z or the evaluations: k synthetic divisions by a linear
//                    factor, each a first-order linear recurrence  q[i] = p[i+1] + a*q[i+1]
//                    evaluated as a suffix scan inside one CTA.
//   poly_eval_range  y_j = P(first_root + j)   (evaluate_polynomial_points, src/util.cpp:186-211)
//
// Coefficients travel as little-endian SCALAR_BYTES records (BytesFromZZ layout, src/util.cpp:31-43)
// and are held in Montgomery form on device.  BN curves have 2-adicity 2, so nothing here uses an NTT.
#pragma once
#if defined(__CUDACC__)
#include <cooperative_groups.h>
#endif
#include "field.cuh"
#include "grid_barrier.cuh"

namespace kzgb200 {

template <class C>
KZG_D Field<typename C::FrP> fr_load(const uint8_t* src, uint64_t idx) {
  using Fr = Field<typename C::FrP>;
  const uint32_t* w = reinterpret_cast<const uint32_t*>(src + idx * C::SCALAR_BYTES);
  Fr k;
#pragma unroll
  for (int i = 0; i < Fr::L; i++) k.v[i] = (i < C::SCALAR_WORDS) ? w[i] : 0u;
  // the Montgomery product reduces any input below 2^(32L) when multiplied by R^2 < r
  return Fr::to_mont(k);
}
template <class C>
KZG_D void fr_store(uint8_t* dst, uint64_t idx, const Field<typename C::FrP>& m) {
  using Fr = Field<typename C::FrP>;
  Fr k = Fr::from_mont(m);
  uint32_t* w = reinterpret_cast<uint32_t*>(dst + idx * C::SCALAR_BYTES);
#pragma unroll
  for (int i = 0; i < C::SCALAR_WORDS; i++) w[i] = k.v[i];
}

// (first_root + j) mod r in Montgomery form; first_root may be negative
template <class C>
__device__ __noinline__ Field<typename C::FrP> fr_from_i64(long long x) {
  using Fr = Field<typename C::FrP>;
  Fr k = Fr::zero();
  unsigned long long mag = x < 0 ? (unsigned long long)(-(x + 1)) + 1ull : (unsigned long long)x;
  k.v[0] = (uint32_t)mag;
  k.v[1] = (uint32_t)(mag >> 32);
  Fr m = Fr::to_mont(k);
  return x < 0 ? Fr::neg(m) : m;
}

template <class Fr>
__device__ __noinline__ Fr fr_pow_u32(Fr base, uint32_t e) {
  Fr r = Fr::one();
  while (e) {
    if (e & 1) r = Fr::mul(r, base);
    base = Fr::sqr(base);
    e >>= 1;
  }
  return r;
}

// bytes -> Montgomery (grid-stride)
template <class C>
__global__ void fr_from_bytes(uint64_t n, const uint8_t* __restrict__ in, Field<typename C::FrP>* __restrict__ out) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    out[i] = fr_load<C>(in, i);
}
template <class C>
__global__ void fr_to_bytes(uint64_t n, const Field<typename C::FrP>* __restrict__ in, uint8_t* __restrict__ out) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    fr_store<C>(out, i, in[i]);
}

// batch of polynomials stored with stride `in_stride`: the first nq coefficients of each -> contiguous records
template <class C>
__global__ void fr_to_bytes_strided(uint64_t nq, uint64_t batch, uint64_t in_stride, const Field<typename C::FrP>* __restrict__ in,
                                    uint8_t* __restrict__ out) {
  const uint64_t total = nq * batch;
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < total; t += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t b = t / nq, i = t - b * nq;
    fr_store<C>(out, t, in[b * in_stride + i]);
  }
}

// ---------------------------------------------------------------- quotient by k consecutive linear factors
// Single CTA.  buf0 holds P (n coefficients, Montgomery); buf1 is scratch of the same size.
// After round j the current polynomial has n-1-j coefficients.  The result (n-k coefficients) ends
// in buf0 if k is even, buf1 if k is odd.
template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
poly_quotient_rounds(uint64_t n, long long first_root, const long long* __restrict__ first_roots, uint64_t k,
                     Field<typename C::FrP>* buf0, Field<typename C::FrP>* buf1) {
  using Fr = Field<typename C::FrP>;
  __shared__ Fr s_val[THREADS];
  __shared__ Fr s_mul[THREADS];
  const int tid = threadIdx.x;
  // batched form: CTA b works on polynomial b (stride n) opened at first_roots[b]
  if (first_roots) first_root = first_roots[blockIdx.x];
  Fr* src = buf0 + (uint64_t)blockIdx.x * n;
  Fr* dst = buf1 + (uint64_t)blockIdx.x * n;
  uint64_t m = n;  // current coefficient count
  for (uint64_t round = 0; round < k; round++) {
    const Fr a = fr_from_i64<C>(first_root + (long long)round);
    // thread t owns source indices [lo, hi), segments ordered by index; carry flows from high to low
    const uint64_t seg = (m + THREADS - 1) / THREADS;
    const uint64_t lo = min(seg * (uint64_t)tid, m), hi = min(lo + seg, m);
    // h = sum_{j in [lo,hi)} src[j] * a^(j-lo)
    Fr h = Fr::zero();
    for (uint64_t j = hi; j > lo; j--) h = Fr::add(Fr::mul(h, a), src[j - 1]);
    // value carried INTO this segment from above:  c_t = sum_{j >= hi} src[j] * a^(j-hi)
    // suffix scan over threads of the affine maps  c -> h_t + a^len_t * c
    s_val[tid] = h;
    s_mul[tid] = fr_pow_u32<Fr>(a, (uint32_t)(hi - lo));
    __syncthreads();
    for (int d = 1; d < THREADS; d <<= 1) {
      Fr v = s_val[tid], mu = s_mul[tid];
      Fr v2 = Fr::zero(), mu2 = Fr::one();
      const bool has = tid + d < THREADS;
      if (has) { v2 = s_val[tid + d]; mu2 = s_mul[tid + d]; }
      __syncthreads();
      if (has) {
        // compose: first apply the upper block (tid+d ...), then this one
        s_val[tid] = Fr::add(v, Fr::mul(mu, v2));
        s_mul[tid] = Fr::mul(mu, mu2);
      }
      __syncthreads();
    }
    // s_val[t] = sum_{j >= lo_t} src[j] a^(j-lo_t); the carry into segment t is s_val[t+1]
    Fr c = (tid + 1 < THREADS) ? s_val[tid + 1] : Fr::zero();
    // q[i] = src[i+1] + a*q[i+1]; with c = q-value "at index hi-1":  q[hi-1] = c? no:
    // define r[j] = sum_{j' >= j} src[j'] a^(j'-j)  (so c = r[hi]); then q[j-1] = r[j] for j >= 1.
    for (uint64_t j = hi; j > lo; j--) {
      // here c == r[j]
      if (j >= 1 && j - 1 < m - 1) dst[j - 1] = c;          // q[j-1] = r[j]   (q has m-1 coefficients)
      c = Fr::add(Fr::mul(c, a), src[j - 1]);               // r[j-1]
    }
    __syncthreads();
    Fr* t = src; src = dst; dst = t;
    m -= 1;
  }
}

// ---------------------------------------------------------------- the same k rounds over the WHOLE GPU
// One polynomial that is too long for one CTA (the reference's large-degree path, benchmark/benchmark.cpp:123-136:
// degrees up to 8 388 608, where a single CTA would walk 2^23 coefficients per root): a cooperative launch, one CTA
// per SM, every thread owns a contiguous segment.  Per root: segment Horner sums, a suffix scan of the affine maps
// c -> h + a^len c first inside each CTA and then (redundantly in every CTA, through shared memory) over the CTA
// aggregates, a grid barrier, the second sweep over the segment that writes the quotient, a grid barrier.
// The barrier is grid_barrier.cuh's generation counter.
template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
poly_quotient_rounds_grid(uint64_t n, long long first_root, uint64_t k, Field<typename C::FrP>* buf0,
                          Field<typename C::FrP>* buf1, Field<typename C::FrP>* agg /* 2 * gridDim */, GridBarrier* bar) {
  using Fr = Field<typename C::FrP>;
  __shared__ Fr s_val[THREADS];
  __shared__ Fr s_mul[THREADS];
  const int tid = threadIdx.x;
  const uint64_t T = (uint64_t)gridDim.x * THREADS, gt = (uint64_t)blockIdx.x * THREADS + tid;
  Fr* src = buf0;
  Fr* dst = buf1;
  uint64_t m = n;
  for (uint64_t round = 0; round < k; round++) {
    const Fr a = fr_from_i64<C>(first_root + (long long)round);
    const uint64_t seg = (m + T - 1) / T;
    const uint64_t lo = min(seg * gt, m), hi = min(lo + seg, m);
    Fr h = Fr::zero();
    for (uint64_t j = hi; j > lo; j--) h = Fr::add(Fr::mul(h, a), src[j - 1]);
    s_val[tid] = h;
    s_mul[tid] = fr_pow_u32<Fr>(a, (uint32_t)(hi - lo));
    __syncthreads();
    for (int d = 1; d < THREADS; d <<= 1) {  // suffix scan inside the CTA (as poly_quotient_rounds)
      Fr v = s_val[tid], mu = s_mul[tid];
      Fr v2 = Fr::zero(), mu2 = Fr::one();
      const bool has = tid + d < THREADS;
      if (has) { v2 = s_val[tid + d]; mu2 = s_mul[tid + d]; }
      __syncthreads();
      if (has) {
        s_val[tid] = Fr::add(v, Fr::mul(mu, v2));
        s_mul[tid] = Fr::mul(mu, mu2);
      }
      __syncthreads();
    }
    // carry into this thread's segment from the REST OF THIS CTA (as if nothing lay above the CTA), and the map of
    // the segments between this thread and the top of the CTA
    const Fr c_local = (tid + 1 < THREADS) ? s_val[tid + 1] : Fr::zero();
    const Fr mul_above = (tid + 1 < THREADS) ? s_mul[tid + 1] : Fr::one();
    if (tid == 0) { agg[2 * blockIdx.x] = s_val[0]; agg[2 * blockIdx.x + 1] = s_mul[0]; }
    grid_sync(bar);
    // carry into this CTA: suffix scan over the aggregates of the CTAs above (gridDim <= THREADS)
    __syncthreads();
    if (tid < (int)gridDim.x) { s_val[tid] = agg[2 * tid]; s_mul[tid] = agg[2 * tid + 1]; }
    else { s_val[tid] = Fr::zero(); s_mul[tid] = Fr::one(); }
    __syncthreads();
    for (int d = 1; d < (int)gridDim.x; d <<= 1) {
      Fr v = s_val[tid], mu = s_mul[tid];
      Fr v2 = Fr::zero(), mu2 = Fr::one();
      const bool has = tid + d < THREADS;
      if (has) { v2 = s_val[tid + d]; mu2 = s_mul[tid + d]; }
      __syncthreads();
      if (has) {
        s_val[tid] = Fr::add(v, Fr::mul(mu, v2));
        s_mul[tid] = Fr::mul(mu, mu2);
      }
      __syncthreads();
    }
    const Fr c_cta = (blockIdx.x + 1 < gridDim.x) ? s_val[blockIdx.x + 1] : Fr::zero();
    __syncthreads();
    Fr c = Fr::add(c_local, Fr::mul(mul_above, c_cta));  // r[hi] of this thread's segment
    for (uint64_t j = hi; j > lo; j--) {
      if (j - 1 < m - 1) dst[j - 1] = c;       // q[j-1] = r[j]   (q has m-1 coefficients)
      c = Fr::add(Fr::mul(c, a), src[j - 1]);  // r[j-1]
    }
    grid_sync(bar);
    Fr* t = src; src = dst; dst = t;
    m -= 1;
  }
}

// ---------------------------------------------------------------- quotient, many roots: systolic pipeline
// The scan kernel above costs one CTA-wide scan per root (~70 us), fine for a handful of opened points
// but not for the multi-proof rows (k = 128 ... 4096).  Here thread j is the synthetic-division stage
// of root first_root + j:  y = x + a_j * carry;  carry = y.  The coefficient stream enters stage 0
// top-down, stage j consumes what stage j-1 produced one step earlier and drops its own last output
// (the remainder), so `t` roots are divided out in m + t steps of ONE modmul each instead of t scans.
// One CTA, T stages per pass; the host chains passes for k > T.  dst receives m - t coefficients.
template <class C, int T, int CHK>
__global__ void __launch_bounds__(T)
poly_quotient_systolic(uint64_t m, uint64_t stride, long long first_root, const long long* __restrict__ first_roots,
                       long long root_shift, int t, const Field<typename C::FrP>* __restrict__ src,
                       Field<typename C::FrP>* __restrict__ dst) {
  using Fr = Field<typename C::FrP>;
  __shared__ Fr link[2][T];
  __shared__ Fr feed[CHK];
  const int j = threadIdx.x;
  if (first_roots) first_root = first_roots[blockIdx.x] + root_shift;  // batched form: CTA b = polynomial b
  src += (uint64_t)blockIdx.x * stride;
  dst += (uint64_t)blockIdx.x * stride;
  const Fr a = fr_from_i64<C>(first_root + (long long)j);
  Fr carry = Fr::zero();
  const uint64_t steps = m + (uint64_t)t;
  for (uint64_t s = 0; s < steps; s++) {
    if (s % CHK == 0) {  // refill the feed of stage 0: src top-down, CHK coefficients at a time
      for (int q = j; q < CHK; q += T) {
        const uint64_t idx = s + (uint64_t)q;
        if (idx < m) feed[q] = src[m - 1 - idx];
      }
      __syncthreads();
    }
    const long long u = (long long)s - j;  // index of this stage's input within its stream
    if (j < t && u >= 0 && (uint64_t)u < m - (uint64_t)j) {
      const Fr x = (j == 0) ? feed[s % CHK] : link[s & 1][j - 1];
      const Fr y = Fr::add(x, Fr::mul(a, carry));
      carry = y;
      if ((uint64_t)u + 1 < m - (uint64_t)j) {  // the last output of a stage is the remainder: dropped
        if (j == t - 1) dst[(m - (uint64_t)t - 1) - (uint64_t)u] = y;
        else link[(s + 1) & 1][j] = y;
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------- evaluation at consecutive points
// One CTA per point: thread t Horner-evaluates its slice, scales by x^lo, CTA tree-sums.
template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
poly_eval_range_kernel(uint64_t n, long long first_root, uint64_t k, const Field<typename C::FrP>* __restrict__ coeffs,
                       Field<typename C::FrP>* __restrict__ ys) {
  using Fr = Field<typename C::FrP>;
  __shared__ Fr s_val[THREADS];
  const int tid = threadIdx.x;
  for (uint64_t pt = blockIdx.x; pt < k; pt += gridDim.x) {
    const Fr x = fr_from_i64<C>(first_root + (long long)pt);
    const uint64_t seg = (n + THREADS - 1) / THREADS;
    const uint64_t lo = min(seg * (uint64_t)tid, n), hi = min(lo + seg, n);
    Fr h = Fr::zero();
    for (uint64_t j = hi; j > lo; j--) h = Fr::add(Fr::mul(h, x), coeffs[j - 1]);
    if (hi > lo) h = Fr::mul(h, fr_pow_u32<Fr>(x, (uint32_t)lo));
    s_val[tid] = h;
    __syncthreads();
    for (int s = THREADS / 2; s > 0; s >>= 1) {
      if (tid < s) s_val[tid] = Fr::add(s_val[tid], s_val[tid + s]);
      __syncthreads();
    }
    if (tid == 0) ys[pt] = s_val[0];
    __syncthreads();
  }
}


// ======================================================================================================
// Interpolation through consecutive integer nodes x_i = first_x + i: replaces
//   polyfit / linear_roots_and_polyfit   src/util.cpp:172-184, 213-248, 269-284   (subproduct tree over NTL mul)
// for the node sets this library's blobs always have (blob::from_bytes / from_string number their chunks
// consecutively, src/blob.cpp:20-48).  Same polynomial, so same canonical coefficients, computed as
//   1. Newton coefficients  c_j = Delta^j y_0 / j! = sum_{i<=j} (y_i / i!) * ((-1)^(j-i) / (j-i)!)
//      — one triangular convolution, a warp per output;
//   2. a product tree in the Newton basis: for a node range split into L | R,
//          P = P_L + Z_L * P_R,   Z = Z_L * Z_R,   Z_L = prod_{t in L} (x - x_t),
//      schoolbook products (no NTT: r - 1 has 2-adicity 2 on the BN curves), a warp per output coefficient,
//      one launch per tree level.  The root holds I (first k coefficients) and Z (k + 1 coefficients).
// k is padded to a power of two with c_j = 0, Z_j = 1 leaves, which changes neither I nor Z.

// inv_fact[t] = 1 / t!  for t < n  (Montgomery), cached by the context.  One CTA: thread u owns a contiguous range of
// integers, a suffix scan of the range products gives (n-1)! and every prod_{j >= hi_u} j, ONE inversion, then each
// thread walks its range downwards: 1/(i-1)! = (1/i!) * i.  (Round 1 ran the 2n products on a single thread: 3 ms for
// n = 4096, the first-call cliff of the multi-proof rows.)
constexpr int INV_FACT_THREADS = 1024;
template <class C>
__global__ void __launch_bounds__(INV_FACT_THREADS)
fr_inv_factorials(uint32_t n, Field<typename C::FrP>* __restrict__ inv_fact) {
  using Fr = Field<typename C::FrP>;
  __shared__ Fr s_prod[INV_FACT_THREADS];
  __shared__ Fr s_inv;
  const int tid = threadIdx.x, T = blockDim.x;
  const uint32_t per = (n + T - 1) / T;
  const uint32_t lo = min(per * (uint32_t)tid, n), hi = min(lo + per, n);
  const Fr one = Fr::one();
  // product of the integers of [lo, hi) (0 counts as 1: 0! = 1)
  Fr p = one, x = fr_from_i64<C>((long long)lo);
  for (uint32_t i = lo; i < hi; i++) {
    if (i > 0) p = Fr::mul(p, x);
    x = Fr::add(x, one);
  }
  s_prod[tid] = p;
  __syncthreads();
  for (int d = 1; d < T; d <<= 1) {  // inclusive suffix scan: s_prod[t] = prod of the ranges t, t+1, ...
    Fr v = s_prod[tid];
    const bool has = tid + d < T;
    Fr v2 = has ? s_prod[tid + d] : one;
    __syncthreads();
    if (has) s_prod[tid] = Fr::mul(v, v2);
    __syncthreads();
  }
  if (tid == 0) s_inv = Fr::inv(s_prod[0]);  // 1 / (n-1)!
  __syncthreads();
  // 1 / (hi-1)! = (1 / (n-1)!) * prod_{j >= hi} j
  Fr inv = Fr::mul(s_inv, tid + 1 < T ? s_prod[tid + 1] : one);
  x = fr_from_i64<C>((long long)hi - 1);
  for (uint32_t i = hi; i-- > lo;) {
    inv_fact[i] = inv;              // 1 / i!
    if (i > 0) inv = Fr::mul(inv, x);  // 1/(i-1)! = (1/i!) * i
    x = Fr::sub(x, one);
  }
}

template <class Fr>
__device__ __forceinline__ Fr warp_sum_field(Fr v) {
  for (int d = 16; d > 0; d >>= 1) {
    Fr o;
#pragma unroll
    for (int i = 0; i < Fr::L; i++) o.v[i] = __shfl_down_sync(0xffffffffu, v.v[i], d);
    v = Fr::add(v, o);
  }
  return v;  // lane 0 holds the sum
}

// blob::from_bytes (src/blob.cpp:19-48) on the device: chunk i of chunk_size bytes, zero-padded, read as a
// little-endian integer (ZZFromBytes) becomes the SB-byte scalar y_i.  chunk_size <= SB - 1 = MAX_CHUNK_BYTES
// (src/kzg.h:31), so the value is below the group order and conv<ZZ_p> is the identity.
static __global__ void blob_encode_chunks(uint64_t n_chunks, int chunk_size, int sb, const uint8_t* __restrict__ bytes,
                                          uint8_t* __restrict__ ys) {
  const uint64_t total = n_chunks * (uint64_t)sb;
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < total; t += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t i = t / (uint64_t)sb;
    const int j = (int)(t % (uint64_t)sb);
    ys[t] = j < chunk_size ? bytes[i * (uint64_t)chunk_size + j] : (uint8_t)0;
  }
}

// u_i = y_i / i!  (y from little-endian bytes)
template <class C>
__global__ void interp_scale_values(uint64_t k, const uint8_t* __restrict__ ys, const Field<typename C::FrP>* __restrict__ inv_fact,
                                    Field<typename C::FrP>* __restrict__ u) {
  using Fr = Field<typename C::FrP>;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < k; i += (uint64_t)gridDim.x * blockDim.x)
    u[i] = Fr::mul(fr_load<C>(ys, i), inv_fact[i]);
}

// leaves of the product tree (level m = 1) for kp = padded k:
//   P0[j] = c_j = sum_{i<=j} u_i * (-1)^(j-i) inv_fact[j-i]   (0 for j >= k)
//   Z0[j] = (x - (first_x + j)) as 2 coefficients  [-(first_x+j), 1]   ([1, 0] for j >= k)
template <class C>
__global__ void __launch_bounds__(128)
interp_leaves(uint64_t k, uint64_t kp, long long first_x, const Field<typename C::FrP>* __restrict__ u,
              const Field<typename C::FrP>* __restrict__ inv_fact, Field<typename C::FrP>* __restrict__ P0,
              Field<typename C::FrP>* __restrict__ Z0, int with_p) {
  using Fr = Field<typename C::FrP>;
  const int lane = threadIdx.x & 31;
  const uint64_t warp = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t j = warp; j < kp; j += nwarps) {
    Fr acc = Fr::zero();
    if (with_p && j < k) {
      for (uint64_t i = lane; i <= j; i += 32) {
        const Fr term = Fr::mul(u[i], inv_fact[j - i]);
        acc = ((j - i) & 1) ? Fr::sub(acc, term) : Fr::add(acc, term);
      }
    }
    acc = warp_sum_field<Fr>(acc);
    if (lane == 0) {
      if (with_p) P0[j] = acc;
      if (j < k) {
        Z0[2 * j] = Fr::neg(fr_from_i64<C>(first_x + (long long)j));
        Z0[2 * j + 1] = Fr::one();
      } else {
        Z0[2 * j] = Fr::one();
        Z0[2 * j + 1] = Fr::zero();
      }
    }
  }
}

// One tree level: nodes of size m (P: m coefficients, Z: m + 1) are merged in pairs into nodes of size 2m.
//   Pout[q][i] = (i < m ? P_L[i] : 0) + sum_t Z_L[t] * P_R[i - t],   i < 2m
//   Zout[q][i] = sum_t Z_L[t] * Z_R[i - t],                           i <= 2m
// A warp per output coefficient, lanes stride over t.
template <class C>
__global__ void __launch_bounds__(128)
interp_tree_level(uint64_t m, uint64_t npairs, const Field<typename C::FrP>* __restrict__ Pin,
                  const Field<typename C::FrP>* __restrict__ Zin, Field<typename C::FrP>* __restrict__ Pout,
                  Field<typename C::FrP>* __restrict__ Zout, int with_p) {
  using Fr = Field<typename C::FrP>;
  const int lane = threadIdx.x & 31;
  const uint64_t warp = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const uint64_t p_outs = with_p ? 2 * m : 0;
  const uint64_t per_pair = p_outs + 2 * m + 1;  // 2m outputs of P (when wanted), 2m + 1 of Z
  for (uint64_t w = warp; w < npairs * per_pair; w += nwarps) {
    const uint64_t q = w / per_pair, o = w - q * per_pair;
    const bool is_z = o >= p_outs;
    const uint64_t i = is_z ? o - p_outs : o;
    const Fr* ZL = Zin + (2 * q) * (m + 1);
    const Fr* R = is_z ? Zin + (2 * q + 1) * (m + 1) : Pin + (2 * q + 1) * m;
    const uint64_t rlen = is_z ? m + 1 : m;
    const uint64_t t_lo = i >= rlen ? i - rlen + 1 : 0, t_hi = i < m ? i : m;  // t in [t_lo, t_hi]
    Fr acc = Fr::zero();
    for (uint64_t t = t_lo + lane; t <= t_hi; t += 32) acc = Fr::add(acc, Fr::mul(ZL[t], R[i - t]));
    acc = warp_sum_field<Fr>(acc);
    if (lane == 0) {
      if (is_z) Zout[q * (2 * m + 1) + i] = acc;
      else Pout[q * (2 * m) + i] = (i < m) ? Fr::add(acc, Pin[(2 * q) * m + i]) : acc;
    }
  }
}

// q = P div Z for a monic Z of degree k (Z = prod (x - x_i), from the product tree above), one CTA:
// schoolbook long division with the k-coefficient active window of the remainder held in registers.
// Coefficient index m of the remainder lives on residue m mod k (thread = residue mod T, slot = residue / T),
// so no value ever moves between threads: at step i (quotient index, downwards) the owner of index i + k holds
// a finished coefficient = q_i, publishes it through shared memory, takes P[i] as its new element, and every
// element m in [i, i + k) subtracts q_i * z_{m-i}.  One barrier and one (EPT independent) modmul per step,
// n - k steps — against (n + 512) steps per pass of 512 roots for the systolic form.
template <class C, int EPT>
__global__ void __launch_bounds__(1024)
poly_long_division(uint64_t n, uint32_t k, const Field<typename C::FrP>* __restrict__ P,
                   const Field<typename C::FrP>* __restrict__ Z, Field<typename C::FrP>* __restrict__ q) {
  using Fr = Field<typename C::FrP>;
  extern __shared__ __align__(16) uint8_t division_smem[];
  Fr* zs = reinterpret_cast<Fr*>(division_smem);  // z_0 .. z_{k-1}
  Fr* qs = zs + k;                                // q_i, double-buffered
  const uint32_t t = threadIdx.x, T = blockDim.x;
  for (uint32_t j = t; j < k; j += T) zs[j] = Z[j];
  Fr r[EPT];
  const uint64_t base = n - k;  // the window starts as the top k coefficients
#pragma unroll
  for (int e = 0; e < EPT; e++) {
    const uint32_t res = t + (uint32_t)e * T;
    r[e] = Fr::zero();
    if (res < k) r[e] = P[base + (uint64_t)((res + k - (uint32_t)(base % k)) % k)];
  }
  __syncthreads();
  uint32_t own = (uint32_t)((n - 1) % k);  // residue of index i + k for i = n - k - 1
  for (int64_t i = (int64_t)(n - k) - 1; i >= 0; i--) {
    if (own % T == t) {
      const uint32_t slot = own / T;
#pragma unroll
      for (int e = 0; e < EPT; e++)
        if ((uint32_t)e == slot) {
          qs[i & 1] = r[e];
          q[i] = r[e];
          r[e] = P[i];
        }
    }
    __syncthreads();
    const Fr qq = qs[i & 1];
#pragma unroll
    for (int e = 0; e < EPT; e++) {
      const uint32_t res = t + (uint32_t)e * T;
      if (res < k) {
        const uint32_t j = res >= own ? res - own : res + k - own;  // m - i for the index m = res (mod k) in the window
        r[e] = Fr::sub(r[e], Fr::mul(qq, zs[j]));
      }
    }
    own = own == 0 ? k - 1 : own - 1;
  }
}

// The same long division on a thread-block CLUSTER of 8 CTAs: the window of k coefficients is spread over
// 8 SMs (one element per thread, residue = cluster-wide thread index), every CTA keeps its own copy of Z in
// shared memory, and the owner of the finished coefficient writes q_i into all 8 CTAs' shared memory
// through DSMEM before the per-step cluster barrier.  A step is then one modmul deep (~0.4 us) plus the
// barrier (~0.2 us) instead of k modmuls queued on one SM's multiplier pipe.
constexpr int DIV_CLUSTER = 8;
template <class C>
__global__ void __cluster_dims__(DIV_CLUSTER, 1, 1) __launch_bounds__(512)
poly_long_division_cluster(uint64_t n, uint32_t k, const Field<typename C::FrP>* __restrict__ P,
                           const Field<typename C::FrP>* __restrict__ Z, Field<typename C::FrP>* __restrict__ q) {
  namespace cg = cooperative_groups;
  using Fr = Field<typename C::FrP>;
  extern __shared__ __align__(16) uint8_t division_smem[];
  Fr* zs = reinterpret_cast<Fr*>(division_smem);
  Fr* qs = zs + k;
  cg::cluster_group cluster = cg::this_cluster();
  const uint32_t t = threadIdx.x, T = blockDim.x;
  const uint32_t res = cluster.block_rank() * T + t;  // the residue (index mod k) this thread owns
  for (uint32_t j = t; j < k; j += T) zs[j] = Z[j];
  Fr* remote_qs[DIV_CLUSTER];
#pragma unroll
  for (int c = 0; c < DIV_CLUSTER; c++) remote_qs[c] = cluster.map_shared_rank(qs, c);
  const uint64_t base = n - k;
  Fr r = Fr::zero(), nxt = Fr::zero();
  if (res < k) {
    r = P[base + (uint64_t)((res + k - (uint32_t)(base % k)) % k)];
    // the element this thread takes over when it first owns the finished coefficient (step i = res mod k),
    // loaded now and refreshed k steps ahead each time, so no global-load latency sits on the step's critical path
    const uint64_t d = ((base - 1) % k + k - res) % k;
    if (base >= 1 + d) nxt = P[base - 1 - d];
  }
  cluster.sync();
  uint32_t own = (uint32_t)((n - 1) % k);
  for (int64_t i = (int64_t)(n - k) - 1; i >= 0; i--) {
    if (own == res) {
#pragma unroll
      for (int c = 0; c < DIV_CLUSTER; c++) remote_qs[c][i & 1] = r;
      q[i] = r;
      r = nxt;
      if (i >= (int64_t)k) nxt = P[i - (int64_t)k];
    }
    cluster.sync();
    if (res < k) {
      const uint32_t j = res >= own ? res - own : res + k - own;
      r = Fr::sub(r, Fr::mul(qs[i & 1], zs[j]));
    }
    own = own == 0 ? k - 1 : own - 1;
  }
  cluster.sync();  // no CTA exits while a peer may still write into its shared memory
}

// ---- quotient by power-series inversion (one polynomial, many roots): q = P div Z with deg Z = k, L = n - k:
//   rev(q) = rev(P) * rev(Z)^-1  mod x^L,   rev(Z)^-1 by Newton doubling  g <- g (2 - f g) mod x^(2m).
// Every step is a truncated schoolbook product, a warp per output coefficient: ~3.2 L^2 modmuls in all, spread
// over the whole GPU in ~3 log2 L launches, instead of L dependent steps on one CTA / one cluster.
// out[i] = sum_t a[t] b[i - t] (i < nout);  two_minus: out = 2 [i == 0] - that.  With a_rev / b_rev the operand
// is read back to front (a[t] := a_base[a_last - t]), which is how rev(P) and rev(Z) are consumed in place, and
// out_rev writes the result back to front (out[nout - 1 - i]).
// Batched form (blockIdx.y = polynomial): a and out advance by a_stride / out_stride coefficients per polynomial
// while b is shared — the last product of the quotient, rev(P_y) * rev(Z)^-1, for every polynomial of a batch that
// is opened at the same points.  With `sel`, polynomial y takes part only if sel[y] == sel_value.
template <class C>
__global__ void __launch_bounds__(128)
poly_mul_trunc(const Field<typename C::FrP>* __restrict__ a, uint64_t na, long long a_last,
               const Field<typename C::FrP>* __restrict__ b, uint64_t nb, long long b_last,
               Field<typename C::FrP>* __restrict__ out, uint64_t nout, int two_minus, int out_rev,
               uint64_t a_stride = 0, uint64_t out_stride = 0, const long long* __restrict__ sel = nullptr,
               long long sel_value = 0) {
  using Fr = Field<typename C::FrP>;
  if (sel && sel[blockIdx.y] != sel_value) return;
  a += blockIdx.y * a_stride;
  out += blockIdx.y * out_stride;
  const int lane = threadIdx.x & 31;
  const uint64_t warp = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t i = warp; i < nout; i += nwarps) {
    const uint64_t t_lo = i >= nb ? i - nb + 1 : 0, t_hi = i < na ? i : na - 1;
    Fr acc = Fr::zero();
    for (uint64_t t = t_lo + lane; t <= t_hi; t += 32) {
      const Fr x = a_last >= 0 ? a[a_last - (long long)t] : a[t];
      const Fr y = b_last >= 0 ? b[b_last - (long long)(i - t)] : b[i - t];
      acc = Fr::add(acc, Fr::mul(x, y));
    }
    acc = warp_sum_field<Fr>(acc);
    if (lane == 0) {
      if (two_minus) {
        acc = Fr::neg(acc);
        if (i == 0) acc = Fr::add(acc, Fr::dbl(Fr::one()));
      }
      out[out_rev ? nout - 1 - i : i] = acc;
    }
  }
}

}  // namespace kzgb200
