// Powers-of-tau batch: [s^i]G1 and [s^i]G2 for i < n, replacing
//   trusted_setup::trusted_setup(int)            src/trusted_setup.cpp:21-74
//   trusted_setup::generate_elements_range       src/trusted_setup.cpp:123-135
// (n independent NTL power() calls, then one PAIR_G1mul and one PAIR_G2mul per index spread over
// std::threads).  Here: s^i by chunked running products, then a fixed-base comb — a table
// T[w][d-1] = d * 2^(16w) * G (d = 1..32768, affine) built once per process and device, 16-bit signed
// digits, one mixed add per window (16 per index for the 254/255-bit orders; round 1 used 8-bit windows = 32
// adds per index and spent 53 ms on 2^20 indices, profiles/r1o_setup.md), one inversion per 8 output points.
// Also: the per-SRS precomputed window table used by the MSM (msm.cuh).
#pragma once
#include "msm.cuh"
#include "msm_team.cuh"
#include "poly.cuh"

namespace kzgb200 {

constexpr int FB_C = 16;                      // fixed-base window bits
constexpr int FB_HALF = 1 << (FB_C - 1);      // 32768 table entries per window (G1 + G2, BLS12-381: 16 x 32768 x 288 B = 151 MB)

template <class C>
KZG_HD constexpr int fb_windows() { return (C::SCALAR_BITS + 1 + FB_C - 1) / FB_C; }

// powers[i] = s^(first_index + i) (canonical little-endian words, NOT Montgomery), i < n.
// Thread t produces indices [t*S, (t+1)*S): s^(first_index + t*S) by square-and-multiply, then S-1 products.
template <class C>
__global__ void setup_powers_kernel(uint64_t n, uint64_t first_index, const uint8_t* __restrict__ s_bytes, int per_thread,
                             uint32_t* __restrict__ powers /* n * FrP::L words */) {
  using Fr = Field<typename C::FrP>;
  const uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  const uint64_t lo = t * (uint64_t)per_thread;
  if (lo >= n) return;
  const uint64_t hi = min(lo + (uint64_t)per_thread, n);
  uint32_t k[Fr::L];
  load_scalar<C>(s_bytes, 0, k);
  Fr s;
#pragma unroll
  for (int i = 0; i < Fr::L; i++) s.v[i] = k[i];
  s = Fr::to_mont(s);
  // s^(first_index + lo), exponent < 2^64
  Fr cur = Fr::one(), base = s;
  for (uint64_t e = first_index + lo; e; e >>= 1) {
    if (e & 1) cur = Fr::mul(cur, base);
    base = Fr::sqr(base);
  }
  for (uint64_t i = lo; i < hi; i++) {
    Fr c = Fr::from_mont(cur);
#pragma unroll
    for (int j = 0; j < Fr::L; j++) powers[i * Fr::L + j] = c.v[j];
    cur = Fr::mul(cur, s);
  }
}

// table[w * FB_HALF + (d-1)] = d * 2^(FB_C w) * G, built in FB_C launches:
//   fixed_base_rows   B_w = 2^(FB_C w) G for every window: one 4-lane team per window runs its chain of doublings
//                     (three product rounds per doubling instead of nine on one thread; the longest chain, 240
//                     doublings, is the latency of the whole build);
//   fixed_base_fill   level j = 0 .. FB_C-2 doubles the filled prefix of every row: (2^j + r) B = 2^j B + r B for
//                     r = 1 .. 2^j, one affine + affine addition and one inversion per new entry (r = 2^j is the
//                     doubling, which the mixed addition's P + P branch takes).
// Round 1 ran a full double-and-add scalar multiplication per entry (~2400 products each); this is ~25.
#if defined(__CUDACC__)
template <class C, class F>
__global__ void fixed_base_rows(Affine<F> gen, Affine<F>* __restrict__ table) {
  constexpr int W = fb_windows<C>();
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> TeamWidth<F>::SHIFT;
  if (w >= W) return;  // whole teams leave together
  XYZZ<F> acc = XYZZ<F>::from_affine(gen);
  for (int i = 0; i < w * FB_C; i++) acc = team_dbl<F>(acc);
  if (Team<F>().lead()) table[(size_t)w * FB_HALF] = acc.to_affine();
}
template <class C, class F>
__global__ void __launch_bounds__(128) fixed_base_fill(int level, Affine<F>* __restrict__ table) {
  constexpr int W = fb_windows<C>();
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t per = 1u << level;
  if (t >= (uint32_t)W * per) return;
  const uint32_t w = t >> level, r = (t & (per - 1u)) + 1u;  // new multiple d = 2^level + r
  Affine<F>* row = table + (size_t)w * FB_HALF;
  XYZZ<F> acc = XYZZ<F>::from_affine(row[per - 1]);
  acc.madd(row[r - 1], false);
  row[per + r - 1] = acc.to_affine();
}
#endif

// sums[i] = powers[i] * G in XYZZ coordinates (normalised afterwards by fixed_base_to_affine: a per-thread
// inversion here is a divergent binary-GCD loop that costs a warp about as much as its 33 mixed additions,
// profiles/r1o_setup.md)
template <class C, class F>
__global__ void __launch_bounds__(128)
fixed_base_mul(uint64_t n, const uint32_t* __restrict__ powers, const Affine<F>* __restrict__ table,
               XYZZ<F>* __restrict__ sums) {
  using Fr = Field<typename C::FrP>;
  constexpr int W = fb_windows<C>();
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t k[Fr::L];
#pragma unroll
  for (int j = 0; j < Fr::L; j++) k[j] = powers[i * Fr::L + j];
  XYZZ<F> acc = XYZZ<F>::inf();
  uint32_t carry = 0;
  for (int w = 0; w < W; w++) {
    uint32_t raw = scalar_bits<Fr::L>(k, w * FB_C, FB_C) + carry;
    int d;
    if (raw > (uint32_t)FB_HALF) { d = (int)raw - (1 << FB_C); carry = 1; }
    else { d = (int)raw; carry = 0; }
    if (d != 0) {
      const int mag = d < 0 ? -d : d;
      const Affine<F> p = table[w * FB_HALF + (mag - 1)];
      acc.madd(p, d < 0);
    }
  }
  sums[i] = acc;
}

// out[i] = sums[i] as Montgomery affine points and as affine bytes (reference octet body layout).
// Thread t normalises the FB_BATCH consecutive sums [t*FB_BATCH, ..) with ONE inversion (Montgomery's trick over
// the ZZZ coordinates: 3 extra products per point); infinity (ZZ = 0) is skipped and written as (0, 0).
constexpr int FB_BATCH = 8;
template <class F>
__global__ void __launch_bounds__(128)
fixed_base_to_affine(uint64_t n, const XYZZ<F>* __restrict__ sums, Affine<F>* __restrict__ out_points,
                     uint8_t* __restrict__ out_bytes, int modbytes) {
  const uint64_t lo = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) * FB_BATCH;
  if (lo >= n) return;
  const int cnt = (int)min((uint64_t)FB_BATCH, n - lo);
  const int nbytes = CoordCount<F>::value * modbytes;
  F pref[FB_BATCH];  // pref[j] = prod_{v<=j} ZZZ_v over the finite sums
  F run = F::one();
  for (int j = 0; j < cnt; j++) {
    const F zzz = sums[lo + j].ZZZ;
    if (!sums[lo + j].ZZ.is_zero()) run = F::mul(run, zzz);
    pref[j] = run;
  }
  F inv = F::inv(run);
  for (int j = cnt - 1; j >= 0; j--) {
    const XYZZ<F> p = sums[lo + j];
    Affine<F> a = Affine<F>::inf();
    if (!p.is_inf()) {
      const F i3 = j > 0 ? F::mul(inv, pref[j - 1]) : inv;  // 1 / ZZZ_j
      inv = F::mul(inv, p.ZZZ);
      const F i1 = F::mul(p.ZZ, i3);             // 1 / Z
      const F i2 = F::sqr(i1);                   // 1 / ZZ
      a = Affine<F>{F::mul(p.X, i2), F::mul(p.Y, i3)};
    }
    if (out_points) out_points[lo + j] = a;
    if (out_bytes) {
      uint8_t* dst = out_bytes + (lo + j) * (uint64_t)nbytes;
      if (p.is_inf()) { for (int b = 0; b < nbytes; b++) dst[b] = 0; }
      else affine_to_bytes(a, dst, modbytes);
    }
  }
}

// The same table with FOUR lanes per point (msm_team.cuh): the W * c doublings of a point are one dependent chain, and
// for the SRS sizes the reference's callers use (a few thousand points) that chain, not the multiplier pipe, is the
// build time — 256 doublings of 9 dependent products each on one thread (12 ms for the G2 table of a 5000-point
// BLS12-381 SRS in round 1).  A team runs a doubling in 3 product rounds.  The normalisation is executed by all four
// lanes on replicated values (same instructions, same warp: no extra issue slots) and written by the first.
#if defined(__CUDACC__)
template <class F, int MAXW>
__global__ void __launch_bounds__(128)
msm_precompute_table_team(uint64_t n, uint64_t stride, int c, int W, const Affine<F>* __restrict__ base,
                          Affine<F>* __restrict__ table_ptr, size_t point_stride) {
  const uint64_t i = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> TeamWidth<F>::SHIFT;
  if (i >= n) return;  // whole teams leave together
  const bool lead = Team<F>().lead();
  struct Strided {
    char* p; size_t st;
    __device__ Affine<F>& operator[](uint64_t k) const { return *reinterpret_cast<Affine<F>*>(p + k * st); }
  } table{reinterpret_cast<char*>(table_ptr), point_stride};
  const Affine<F> p0 = base[i];
  if (lead) table[i] = p0;
  if (p0.is_inf()) {
    if (lead) for (int w = 1; w < W; w++) table[(uint64_t)w * stride + i] = p0;
    return;
  }
  XYZZ<F> cur = XYZZ<F>::from_affine(p0);
  XYZZ<F> pts[MAXW];
  F pref[MAXW];
  F run = F::one();
  for (int w = 1; w < W; w++) {
    for (int k = 0; k < c; k++) cur = team_dbl<F>(cur);
    pts[w] = cur;
    run = F::mul(run, cur.is_inf() ? F::one() : cur.ZZZ);
    pref[w] = run;
  }
  F inv = F::inv(run);
  for (int w = W - 1; w >= 1; w--) {
    const F before = (w > 1) ? pref[w - 1] : F::one();
    if (pts[w].is_inf()) { if (lead) table[(uint64_t)w * stride + i] = Affine<F>::inf(); continue; }
    const F i3 = F::mul(inv, before);
    inv = F::mul(inv, pts[w].ZZZ);
    const F i1 = F::mul(pts[w].ZZ, i3);
    const F i2 = F::sqr(i1);
    const Affine<F> out{F::mul(pts[w].X, i2), F::mul(pts[w].Y, i3)};
    if (lead) table[(uint64_t)w * stride + i] = out;
  }
}
#endif

// MSM window table:  table[w * stride + i] = 2^(c*w) * base[i],  w < W  (row 0 = the SRS itself).
// One thread per point; the W projective values are normalised with ONE inversion (Montgomery's
// trick over the thread's own W values), which is why the scratch lives in local memory.
template <class F, int MAXW>
__global__ void __launch_bounds__(128)
msm_precompute_table(uint64_t n, uint64_t stride, int c, int W, const Affine<F>* __restrict__ base,
                     Affine<F>* __restrict__ table_ptr, size_t point_stride) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  // entries lie point_stride bytes apart (sizeof(Affine<F>) normally; 128 in the padded-table experiment)
  struct Strided {
    char* p; size_t st;
    __device__ Affine<F>& operator[](uint64_t k) const { return *reinterpret_cast<Affine<F>*>(p + k * st); }
  } table{reinterpret_cast<char*>(table_ptr), point_stride};
  const Affine<F> p0 = base[i];
  table[i] = p0;
  if (p0.is_inf()) {
    for (int w = 1; w < W; w++) table[(uint64_t)w * stride + i] = p0;
    return;
  }
  XYZZ<F> cur = XYZZ<F>::from_affine(p0);
  XYZZ<F> pts[MAXW];
  F pref[MAXW];  // pref[w] = prod_{v<=w} ZZZ_v
  F run = F::one();
  for (int w = 1; w < W; w++) {
    for (int k = 0; k < c; k++) cur.dbl();
    pts[w] = cur;                      // infinity only for points outside the prime-order subgroup
    run = F::mul(run, cur.is_inf() ? F::one() : cur.ZZZ);
    pref[w] = run;
  }
  F inv = F::inv(run);
  for (int w = W - 1; w >= 1; w--) {
    const F before = (w > 1) ? pref[w - 1] : F::one();
    if (pts[w].is_inf()) { table[(uint64_t)w * stride + i] = Affine<F>::inf(); continue; }
    const F i3 = F::mul(inv, before);          // 1 / ZZZ_w
    inv = F::mul(inv, pts[w].ZZZ);
    const F i1 = F::mul(pts[w].ZZ, i3);
    const F i2 = F::sqr(i1);
    table[(uint64_t)w * stride + i] = Affine<F>{F::mul(pts[w].X, i2), F::mul(pts[w].Y, i3)};
  }
}

}  // namespace kzgb200
