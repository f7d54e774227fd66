// The reduction tail of a G1 MSM in HOMOGENEOUS PROJECTIVE coordinates on 8-lane teams.
//
// Measured (kzg_gpu_microbench kinds 4..7, profiles/r2g_latency.json): one dependent Fp Montgomery product on a warp
// that has its SM sub-partition to itself takes 1.05 us, a 4-lane XYZZ addition (msm_team.cuh: 14 products, FIVE
// dependent rounds) 6.7 us, a grid barrier 2.3 us — a step of msm_tail costs ~10.5 us and a 2^12-point MSM is ~40
// such steps deep.  The depth in product ROUNDS is what can still be cut: the complete addition formulas of Renes,
// Costello and Batina for y^2 = x^3 + b (a = 0; "Complete addition formulas for prime order elliptic curves",
// Algorithms 7 and 9) need 12 products in only TWO rounds of six independent ones, and a doubling 8 products in two
// rounds of four, because the multiplications by 3b are a few additions when b is small (b = 5, 2, 4 on the three
// curves of this library).  One operation per 8 adjacent lanes: lanes 0-5 take the six products of a round, the
// linear steps run replicated.  The formulas are complete for curves of odd order (all three), so infinity (0 : 1 : 0),
// P + P and P + (-P) need no branches.
//
// msm_tail_proj is msm_tail (msm_tail.cuh) with these operations for every step that is narrow enough for the
// 8-lane teams; the fold of the task partials (in place on the XYZZ partials that the wide kernels also read) and
// levels of more than 4 rounds of teams keep XYZZ, converted on first use (X ZZZ : Y ZZ : ZZ ZZZ, one round).
// G2 (the twist constants are tiny too: b' = 5 + 5i, 1 - i, 4 + 4i): with an Fp2 product done whole by one lane (three
// dependent Fp products) the 8-lane form brought nothing (profiles/r2k_probe_g2.json).  ProjTeam<Fp2> therefore spreads
// one operation over the WHOLE WARP: 8 roles x 4 lanes, a role's Fp2 product split into its three Karatsuba Fp
// products (as the 16-lane XYZZ teams of msm_team.cuh), so a G2 addition is also two Fp products deep.
#pragma once
#include "msm_tail.cuh"

#if defined(__CUDACC__)
namespace kzgb200 {

template <class F>
struct Proj {
  F X, Y, Z;
  __device__ __forceinline__ bool is_inf() const { return Z.is_zero(); }
};

// The lanes that share one projective operation.  G1: 8 lanes, one product of the round per lane (roles 0..5 used).
// G2: the whole warp = 8 roles x 4 parts; part 0 / 1 / 2 compute a a' / b b' / (a + b)(a' + b') of the role's Fp2 product.
template <class F> struct ProjTeam;
template <class P>
struct ProjTeam<Field<P>> {
  using F = Field<P>;
  using Piece = F;
  static constexpr int LANES = 8;
  int role, base;
  unsigned mask;
  __device__ __forceinline__ ProjTeam() {
    const int lane = threadIdx.x & 31;
    role = lane & 7;
    base = lane & ~7;
    mask = 0xFFu << base;
  }
  __device__ __forceinline__ Piece piece(const F& u, const F& v) const { return F::mul(u, v); }
  __device__ __forceinline__ F product(const Piece& m, int r) const { return team_shfl(mask, m, base + r); }
};
template <class P>
struct ProjTeam<Field2<P>> {
  using F = Field2<P>;
  using Base = Field<P>;
  using Piece = Base;
  static constexpr int LANES = 32;
  int role, part;
  static constexpr unsigned mask = 0xffffffffu;
  __device__ __forceinline__ ProjTeam() {
    const int lane = threadIdx.x & 31;
    role = lane >> 2;
    part = lane & 3;
  }
  // Karatsuba recombination by the role's own lanes (see team_piece in msm_team.cuh): part 0 ends up with the real,
  // part 1 with the imaginary component of the role's product
  __device__ __forceinline__ Piece piece(const F& u, const F& v) const {
    const Base x = part == 0 ? u.a : part == 1 ? u.b : Base::add(u.a, u.b);
    const Base y = part == 0 ? v.a : part == 1 ? v.b : Base::add(v.a, v.b);
    const Base m = Base::mul(x, y);
    const int lane = threadIdx.x & 31;
    const Base other = team_shfl(mask, m, lane ^ 1), third = team_shfl(mask, m, (lane & ~3) | 2);
    return Base::sub(Base::sub(part == 0 ? m : third, other), part == 0 ? Base::zero() : m);
  }
  __device__ __forceinline__ F product(const Piece& m, int r) const {
    return F{team_shfl(mask, m, 4 * r), team_shfl(mask, m, 4 * r + 1)};
  }
};

// 3b of the curve equation as small integers: 3b for G1 (re), 3b' = re + im i for the twist (G2); all three curve
// configurations have tiny constants here (b = 5, 2, 4; b' = 5 + 5i, 1 - i, 4 + 4i), so the products are additions.
struct B3 {
  int re, im;
};

// k * x for a small non-zero k by doubling and adding
template <class F>
__device__ __forceinline__ F mul_small_pos(const F& x, int k) {
  F r = x;
  int bit = 31 - __clz(k);
  for (bit--; bit >= 0; bit--) {
    r = F::dbl(r);
    if ((k >> bit) & 1) r = F::add(r, x);
  }
  return r;
}
template <class P>
__device__ __forceinline__ Field<P> mul_small(const Field<P>& x, B3 k) {
  return mul_small_pos(x, k.re);
}
// (a + b i)(re + im i) = (a re - b im) + (a im + b re) i
template <class P>
__device__ __forceinline__ Field2<P> mul_small(const Field2<P>& x, B3 k) {
  using Base = Field<P>;
  const int are = k.re < 0 ? -k.re : k.re, aim = k.im < 0 ? -k.im : k.im;
  Base a_re = mul_small_pos(x.a, are), b_re = mul_small_pos(x.b, are);
  if (k.re < 0) { a_re = Base::neg(a_re); b_re = Base::neg(b_re); }
  if (k.im == 0) return Field2<P>{a_re, b_re};
  Base a_im = mul_small_pos(x.a, aim), b_im = mul_small_pos(x.b, aim);
  if (k.im < 0) { a_im = Base::neg(a_im); b_im = Base::neg(b_im); }
  return Field2<P>{Base::sub(a_re, b_im), Base::add(a_im, b_re)};
}

// a + b, complete (RCB Algorithm 7 with a = 0, b3 = 3b).  a, b and the result are replicated on the 8 lanes.
template <class F>
__device__ __noinline__ Proj<F> oct_add(const Proj<F>& a, const Proj<F>& b, B3 b3) {
  const ProjTeam<F> tm;
  const int r = tm.role;
  // round 1:  X1X2 | Y1Y2 | Z1Z2 | (X1+Y1)(X2+Y2) | (Y1+Z1)(Y2+Z2) | (X1+Z1)(X2+Z2)
  F u = r == 0 ? a.X : r == 1 ? a.Y : a.Z, v = r == 0 ? b.X : r == 1 ? b.Y : b.Z;
  if (r >= 3) {  // (uniform inside a team only per lane: both sides are cheap, the compiler predicates them)
    u = F::add(r == 4 ? a.Y : a.X, r == 3 ? a.Y : a.Z);
    v = F::add(r == 4 ? b.Y : b.X, r == 3 ? b.Y : b.Z);
  }
  const auto m1 = tm.piece(u, v);
  const F t0 = tm.product(m1, 0), t1o = tm.product(m1, 1), t2o = tm.product(m1, 2);
  F t3 = tm.product(m1, 3), t4 = tm.product(m1, 4), y3 = tm.product(m1, 5);
  t3 = F::sub(F::sub(t3, t0), t1o);                // X1Y2 + X2Y1
  t4 = F::sub(F::sub(t4, t1o), t2o);               // Y1Z2 + Y2Z1
  y3 = mul_small(F::sub(F::sub(y3, t0), t2o), b3); // b3 (X1Z2 + X2Z1)
  const F x3 = F::add(F::dbl(t0), t0);             // 3 X1X2
  const F t2 = mul_small(t2o, b3);                 // b3 Z1Z2
  const F z3 = F::add(t1o, t2), t1 = F::sub(t1o, t2);
  // round 2:  t4 y3 | t3 t1 | y3 x3 | t1 z3 | x3 t3 | z3 t4
  const F p = r == 0 ? t4 : r == 1 ? t3 : r == 2 ? y3 : r == 3 ? t1 : r == 4 ? x3 : z3;
  const F q = r == 0 ? y3 : r == 1 ? t1 : r == 2 ? x3 : r == 3 ? z3 : r == 4 ? t3 : t4;
  const auto m2 = tm.piece(p, q);
  Proj<F> o;
  o.X = F::sub(tm.product(m2, 1), tm.product(m2, 0));
  o.Y = F::add(tm.product(m2, 3), tm.product(m2, 2));
  o.Z = F::add(tm.product(m2, 5), tm.product(m2, 4));
  return o;
}

// 2 a, complete (RCB Algorithm 9 with a = 0)
template <class F>
__device__ __noinline__ Proj<F> oct_dbl(const Proj<F>& a, B3 b3) {
  const ProjTeam<F> tm;
  const int r = tm.role;
  // round 1:  Y Y | Y Z | Z Z | X Y
  const auto m1 = tm.piece(r == 2 ? a.Z : r >= 3 ? a.X : a.Y, r == 0 || r >= 3 ? a.Y : a.Z);
  const F t0 = tm.product(m1, 0), t1 = tm.product(m1, 1), xy = tm.product(m1, 3);
  const F t2 = mul_small(tm.product(m1, 2), b3);  // b3 Z^2
  const F z8 = F::dbl(F::dbl(F::dbl(t0)));                    // 8 Y^2
  const F y3 = F::add(t0, t2);
  const F t0m = F::sub(t0, F::add(F::dbl(t2), t2));           // Y^2 - 3 b3 Z^2
  // round 2:  t2 z8 | t1 z8 | t0m y3 | t0m xy
  const auto m2 = tm.piece(r == 0 ? t2 : r == 1 ? t1 : t0m, r <= 1 ? z8 : r == 2 ? y3 : xy);
  Proj<F> o;
  o.Y = F::add(tm.product(m2, 0), tm.product(m2, 2));
  o.Z = tm.product(m2, 1);
  o.X = F::dbl(tm.product(m2, 3));
  return o;
}

// two XYZZ values -> projective in ONE round: (X ZZZ : Y ZZ : ZZ ZZZ); XYZZ infinity (ZZ = 0) -> (0 : 1 : 0)
template <class F>
__device__ __noinline__ void oct_from_xyzz2(const XYZZ<F>& a, const XYZZ<F>& b, Proj<F>& pa, Proj<F>& pb) {
  const ProjTeam<F> tm;
  const int r = tm.role;
  const bool second = r >= 3 && r < 6;
  const XYZZ<F>& s = second ? b : a;
  const int k = second ? r - 3 : (r < 3 ? r : 0);
  const auto m = tm.piece(k == 0 ? s.X : k == 1 ? s.Y : s.ZZ, k == 1 ? s.ZZ : s.ZZZ);
  pa.X = tm.product(m, 0); pa.Y = tm.product(m, 1); pa.Z = tm.product(m, 2);
  pb.X = tm.product(m, 3); pb.Y = tm.product(m, 4); pb.Z = tm.product(m, 5);
  if (a.is_inf()) { pa.X = F::zero(); pa.Y = F::one(); pa.Z = F::zero(); }
  if (b.is_inf()) { pb.X = F::zero(); pb.Y = F::one(); pb.Z = F::zero(); }
}

// projective -> XYZZ (X Z : Y Z^2 : Z^2 : Z^3) for the callers that exchange projective partial sums as XYZZ
template <class F>
__device__ __noinline__ XYZZ<F> oct_to_xyzz(const Proj<F>& a) {
  const ProjTeam<F> tm;
  const int r = tm.role;
  const auto m1 = tm.piece(r == 0 ? a.X : a.Z, a.Z);  // X Z | Z^2
  const F xz = tm.product(m1, 0), zz = tm.product(m1, 1);
  const auto m2 = tm.piece(r == 0 ? a.Y : a.Z, zz);   // Y Z^2 | Z^3
  XYZZ<F> o;
  o.X = xz;
  o.Y = tm.product(m2, 0);
  o.ZZ = zz;
  o.ZZZ = tm.product(m2, 1);
  if (a.is_inf()) o = XYZZ<F>::inf();
  return o;
}

template <class F>
__device__ Proj<F> oct_small_mul(const Proj<F>& p, uint32_t k, B3 b3) {
  if (k == 0) return Proj<F>{F::zero(), F::one(), F::zero()};
  Proj<F> m = p;  // the top bit
  for (int bit = top_bit_u32(k) - 1; bit >= 0; bit--) {
    m = oct_dbl<F>(m, b3);
    if ((k >> bit) & 1) m = oct_add<F>(m, p, b3);
  }
  return m;
}

// an array that holds either XYZZ values (written by the wide kernels / thread-mode levels) or projective ones
template <class F>
struct MixedArray {
  XYZZ<F>* p;
  bool xyzz;
  __device__ __forceinline__ Proj<F>* proj() const { return reinterpret_cast<Proj<F>*>(p); }
};

// elements i and j of `in` as projective values (one conversion round when the array still holds XYZZ)
template <class F>
__device__ __forceinline__ void oct_load2(const MixedArray<F>& in, uint64_t i, uint64_t j, Proj<F>& a, Proj<F>& b) {
  if (in.xyzz) {
    const XYZZ<F> x = in.p[i], y = in.p[j];
    oct_from_xyzz2<F>(x, y, a, b);
  } else {
    a = in.proj()[i];
    b = in.proj()[j];
  }
}

struct OctLane {
  uint64_t oct, nocts;
  bool lead;
  int lanes;
  template <class F>
  __device__ __forceinline__ OctLane(const TailLane<F>& tl) {
    constexpr int LANES = ProjTeam<F>::LANES;
    lanes = LANES;
    oct = tl.vwarp * (32 / LANES) + (tl.lane / LANES);
    nocts = tl.nvwarps * (32 / LANES);
    lead = (tl.lane % LANES) == 0;
  }
  // a level runs on the projective teams while that takes at most four rounds
  __device__ __forceinline__ bool fits(uint64_t items) const { return items <= 4 * nocts; }
};

// one level of the two pairwise trees; switches an array to projective the first time its level fits the teams
template <class F>
__device__ __forceinline__ void tailp_pair_level(const TailLane<F>& tl, const OctLane& ol, MixedArray<F>& in0, MixedArray<F>& out0,
                                                 uint64_t n0, MixedArray<F>& in1, MixedArray<F>& out1, uint64_t n1, B3 b3) {
  const bool octs = ol.fits(n0 + n1) || (n0 && !in0.xyzz) || (n1 && !in1.xyzz);
  if (octs) {
    for (uint64_t t = ol.oct; t < n0 + n1; t += ol.nocts) {
      const bool second = t >= n0;
      const uint64_t i = second ? t - n0 : t;
      Proj<F> a, b;
      oct_load2<F>(second ? in1 : in0, 2 * i, 2 * i + 1, a, b);
      const Proj<F> s = oct_add<F>(a, b, b3);
      if (ol.lead) (second ? out1 : out0).proj()[i] = s;
    }
    if (n0) out0.xyzz = false;
    if (n1) out1.xyzz = false;
  } else {  // too wide: XYZZ, in the form msm_tail would pick for this width
    tail_pair_level<F>(tl, in0.p, out0.p, n0, in1.p, out1.p, n1);
    if (n0) out0.xyzz = true;
    if (n1) out1.xyzz = true;
  }
}

template <class F>
__device__ __forceinline__ void tailp_trees(const TailLane<F>& tl, const OctLane& ol, GridBarrier* bar, MixedArray<F>& in0,
                                            MixedArray<F>& out0, uint64_t sums0, uint64_t& len0, MixedArray<F>& in1,
                                            MixedArray<F>& out1, uint64_t sums1, uint64_t& len1, B3 b3) {
  while (len0 > 1 || len1 > 1) {
    tailp_pair_level<F>(tl, ol, in0, out0, len0 > 1 ? sums0 * (len0 / 2) : 0, in1, out1, len1 > 1 ? sums1 * (len1 / 2) : 0, b3);
    if (len0 > 1) { MixedArray<F> t = in0; in0 = out0; out0 = t; len0 /= 2; }
    if (len1 > 1) { MixedArray<F> t = in1; in1 = out1; out1 = t; len1 /= 2; }
    grid_sync(bar);
  }
}

template <class F>
__global__ void __launch_bounds__(TAIL_THREADS, 1) msm_tail_proj(TailArgs<F> a, B3 b3) {
  const TailLane<F> tl;
  const OctLane ol(tl);
  // arrays produced before this launch (wide kernels) hold XYZZ values
  MixedArray<F> rin{a.rin, true}, rout{a.rout, true}, cin{a.cin, true}, cout{a.cout, true};
  bool group_xyzz = true;
  uint64_t r_len = a.r_len, c_len = a.c_len;

  if (a.first_stage <= TAIL_FOLD && a.last_stage >= TAIL_FOLD) {
    // in place on the XYZZ task partials (the wide kernels read them too): XYZZ teams as in msm_tail
    tail_fold_xyzz<F>(tl, a);
  }
  if (a.first_stage <= TAIL_BUCKET_PAIRS && a.last_stage >= TAIL_BUCKET_PAIRS) {
    const uint64_t Lr = a.Cn >= 2 ? a.Cn / 2 : 1, Lc = a.R >= 2 ? a.R / 2 : 1;
    const BucketJob<F> jr{rin.p, a.R, a.Cn, a.Cn, 1, a.groups * a.R * Lr};
    const BucketJob<F> jc{cin.p, a.Cn, a.R, 1, a.Cn, a.groups * a.Cn * Lc};
    const uint64_t total = jr.total + jc.total;
    if (ol.fits(total)) {
      for (uint64_t t0 = ol.oct; t0 < total; t0 += ol.nocts) {
        const bool second = t0 >= jr.total;
        const BucketJob<F>& jb = second ? jc : jr;
        const uint64_t t = second ? t0 - jr.total : t0;
        const uint64_t P = jb.len >= 2 ? jb.len / 2 : 1;
        const uint64_t p = t % P, gs = t / P;
        const uint64_t s = gs % jb.nsums, g = gs / jb.nsums;
        const uint64_t b0 = g * a.nb + s * jb.sstride + 2 * p * jb.jstride;
        const XYZZ<F> x = bucket_value<F>(a.task_first, a.partial, b0);
        const XYZZ<F> y = jb.len >= 2 ? bucket_value<F>(a.task_first, a.partial, b0 + jb.jstride) : XYZZ<F>::inf();
        Proj<F> pa, pb;
        oct_from_xyzz2<F>(x, y, pa, pb);
        const Proj<F> acc = oct_add<F>(pa, pb, b3);
        if (ol.lead) (second ? cin : rin).proj()[t] = acc;
      }
      rin.xyzz = cin.xyzz = false;
    } else {
      tail_bucket_pairs_xyzz<F>(tl, a, jr, jc);
    }
    r_len = Lr;
    c_len = Lc;
    grid_sync(a.barrier);
  }
  if (a.first_stage <= TAIL_TREES1 && a.last_stage >= TAIL_TREES1)
    tailp_trees<F>(tl, ol, a.barrier, rin, rout, a.groups * a.R, r_len, cin, cout, a.groups * a.Cn, c_len, b3);
  if (a.first_stage <= TAIL_WEIGHTED && a.last_stage >= TAIL_WEIGHTED) {
    const uint64_t Pr = a.R >= WCHUNK ? a.R / WCHUNK : 1, Pc = a.Cn >= WCHUNK ? a.Cn / WCHUNK : 1;
    const uint64_t total_r = a.groups * Pr, total_c = a.groups * Pc, total = total_r + total_c;
    if (ol.fits(total) || !rin.xyzz || !cin.xyzz) {
      static_assert(WCHUNK == 2, "the projective weighted step is written for chunks of two");
      for (uint64_t t0 = ol.oct; t0 < total; t0 += ol.nocts) {
        const bool second = t0 >= total_r;
        const uint64_t t = second ? t0 - total_r : t0;
        const uint64_t m = second ? a.Cn : a.R;
        const MixedArray<F>& in = second ? cin : rin;
        const uint32_t woff = second ? 1u : 0u;
        const uint64_t P = m >= WCHUNK ? m / WCHUNK : 1;
        const uint64_t p = t % P, g = t / P;
        const uint64_t j0i = p * WCHUNK;
        // chunk {j0, j0 + 1} (or {j0} alone when m == 1): acc = v0 + 2 v1, run = v0 + v1 (msm_weighted_chunk_thread)
        Proj<F> v0, v1;
        oct_load2<F>(in, g * m + j0i, g * m + (m >= 2 ? j0i + 1 : j0i), v0, v1);
        Proj<F> run = v0, acc = v0;
        if (m >= 2) {
          run = oct_add<F>(v1, v0, b3);
          acc = oct_add<F>(v1, run, b3);
        }
        const uint32_t k = (uint32_t)j0i + woff;  // weights wanted: j + woff = (j - j0 + 1) + (k - 1)
        if (k >= 1) acc = oct_add<F>(acc, oct_small_mul<F>(run, k - 1, b3), b3);
        else {
          run.Y = F::neg(run.Y);
          acc = oct_add<F>(acc, run, b3);
        }
        if (ol.lead) (second ? cout : rout).proj()[t] = acc;
      }
      rout.xyzz = cout.xyzz = false;
    } else {
      const WeightedJob<F> wr{rin.p, rout.p, a.R, total_r, 0u};
      const WeightedJob<F> wc{cin.p, cout.p, a.Cn, total_c, 1u};
      tail_weighted_xyzz<F>(tl, wr, wc);
      rout.xyzz = cout.xyzz = true;
    }
    { MixedArray<F> t = rin; rin = rout; rout = t; }
    { MixedArray<F> t = cin; cin = cout; cout = t; }
    r_len = Pr;
    c_len = Pc;
    grid_sync(a.barrier);
  }
  if (a.first_stage <= TAIL_TREES2 && a.last_stage >= TAIL_TREES2)
    tailp_trees<F>(tl, ol, a.barrier, rin, rout, a.groups, r_len, cin, cout, a.groups, c_len, b3);
  if (a.first_stage <= TAIL_COMBINE && a.last_stage >= TAIL_COMBINE) {
    // group_out[g] = 2^h * T1[g] + T2[g]
    if (ol.fits(a.groups) || !rin.xyzz || !cin.xyzz) {
      Proj<F>* gout = reinterpret_cast<Proj<F>*>(a.group_out);
      for (uint64_t g = ol.oct; g < a.groups; g += ol.nocts) {
        Proj<F> acc, b;
        if (rin.xyzz == cin.xyzz && rin.xyzz) {
          const XYZZ<F> x = rin.p[g], y = cin.p[g];
          oct_from_xyzz2<F>(x, y, acc, b);
        } else {
          Proj<F> dummy;
          if (rin.xyzz) { const XYZZ<F> x = rin.p[g]; oct_from_xyzz2<F>(x, x, acc, dummy); } else acc = rin.proj()[g];
          if (cin.xyzz) { const XYZZ<F> y = cin.p[g]; oct_from_xyzz2<F>(y, y, b, dummy); } else b = cin.proj()[g];
        }
        for (int k = 0; k < a.h; k++) acc = oct_dbl<F>(acc, b3);
        acc = oct_add<F>(acc, b, b3);
        if (a.last_stage >= TAIL_FINISH) {
          if (ol.lead) gout[g] = acc;
        } else {  // a separate msm_finish launch follows (large batches): it reads XYZZ
          const XYZZ<F> x = oct_to_xyzz<F>(acc);
          if (ol.lead) a.group_out[g] = x;
        }
      }
      group_xyzz = a.last_stage < TAIL_FINISH;
    } else {
      tail_combine_xyzz<F>(tl, a, rin.p, cin.p);
    }
    if (a.last_stage >= TAIL_FINISH) grid_sync(a.barrier);
  }
  if (a.first_stage <= TAIL_FINISH && a.last_stage >= TAIL_FINISH) {
    // one MSM per WARP (its first team): the inversion is a data-dependent loop
    if (tl.lane < ProjTeam<F>::LANES) {
      const Proj<F>* gp = reinterpret_cast<const Proj<F>*>(a.group_out);
      for (uint64_t m = tl.vwarp; m < a.batch; m += tl.nvwarps) {
        auto load = [&](uint64_t idx) -> Proj<F> {
          if (!group_xyzz) return gp[idx];
          Proj<F> v, dummy;
          const XYZZ<F> x = a.group_out[idx];
          oct_from_xyzz2<F>(x, x, v, dummy);
          return v;
        };
        Proj<F> acc = load(m * a.groups_per_msm + (a.groups_per_msm - 1));
        for (int g = a.groups_per_msm - 2; g >= 0; g--) {
          for (int k = 0; k < a.c; k++) acc = oct_dbl<F>(acc, b3);
          const Proj<F> v = load(m * a.groups_per_msm + g);
          acc = oct_add<F>(acc, v, b3);
        }
        const bool inf = acc.is_inf();
        if (a.out_xyzz) {
          const XYZZ<F> x = oct_to_xyzz<F>(acc);
          if (tl.lane == 0) a.out_xyzz[m] = x;
        }
        if (tl.lane == 0) {
          const int nbytes = CoordCount<F>::value * a.modbytes;
          if (!a.want_affine) {
            a.is_inf[m] = inf ? 1 : 0;
          } else if (inf) {
            for (int j = 0; j < nbytes; j++) a.out_bytes[m * nbytes + j] = 0;
            a.is_inf[m] = 1;
          } else {
            const F zi = F::inv(acc.Z);
            const Affine<F> aff{F::mul(acc.X, zi), F::mul(acc.Y, zi)};
            affine_to_bytes(aff, a.out_bytes + m * nbytes, a.modbytes);
            a.is_inf[m] = 0;
          }
        }
      }
    }
  }
}

// kzg_gpu_microbench kinds 10..13: one dependent chain of G2 additions on 16-lane XYZZ teams (10), 4-lane XYZZ teams (11),
// whole-warp projective teams (12), and of G1 additions on 8-lane projective teams (13), msm_tail configuration
template <class F1, class F2>
__global__ void __launch_bounds__(TAIL_THREADS, 1) bench_latency_chain_g2(int kind, int iters, uint32_t seed, B3 b3, uint32_t* sink) {
  F1 t = F1::one();
  t.v[0] ^= (seed & 0xffu) + 3u;
  const F1 u = F1::mul(t, F1::r2()), v = F1::sqr(u), w = F1::mul(u, v);
  uint32_t o = 0;
  if (kind == 13) {
    Proj<F1> acc{u, v, w}, p{w, u, v};
    for (int it = 0; it < iters; it++) acc = oct_add<F1>(acc, p, b3);
    o = acc.X.v[0] ^ acc.Z.v[1];
  } else if (kind == 12) {
    Proj<F2> acc{F2{u, v}, F2{v, w}, F2{w, u}}, p{F2{w, v}, F2{u, w}, F2{v, u}};
    for (int it = 0; it < iters; it++) acc = oct_add<F2>(acc, p, b3);
    o = acc.X.a.v[0] ^ acc.Z.b.v[1];
  } else {
    XYZZ<F2> acc, p;
    acc.X = F2{u, v}; acc.Y = F2{v, w}; acc.ZZ = F2{w, u}; acc.ZZZ = F2{u, w};
    p.X = F2{w, v}; p.Y = F2{u, u}; p.ZZ = F2{v, u}; p.ZZZ = F2{w, w};
    for (int it = 0; it < iters; it++) {
      if (kind == 10) acc = team_add<F2, 16>(acc, p);
      else acc = team_add<F2, 4>(acc, p);
    }
    o = acc.X.a.v[0] ^ acc.ZZZ.b.v[1];
  }
  if (o == 0x12345678u) sink[0] = o;
}

}  // namespace kzgb200
#endif
