// The latency-bound tail of the bucket reduction on lane teams (4 lanes for G1, 16 for G2: see TeamWidth).
//
// One thread's XYZZ addition is a chain of 14 dependent Montgomery products (~19 us on B200: every
// product is itself a serialised carry chain).  Once a reduction level has fewer items than the machine
// has warp slots, that chain — not the multiplier pipe — is the cost of the level.  Here four adjacent
// lanes share one group operation: the independent products of each formula step run on different
// lanes and the operands travel by warp shuffles, so an addition is 5 product rounds deep instead of
// 14 and a doubling 3 instead of 9.  It spends ~40 % more multiplier work per operation, so the host
// (ctx.cuh) picks these kernels only for levels of at most TEAM_MAX_ITEMS items (measured cross-over on B200); wide levels keep
// the one-thread-per-item kernels of msm.cuh.  Point values are replicated on the four lanes, which
// keeps every branch uniform inside a team; shuffles name only the team's own lanes in their mask.
#pragma once
#include "msm.cuh"

#if defined(__CUDACC__)
namespace kzgb200 {

constexpr uint64_t TEAM_MAX_ITEMS = 100000;

template <class P>
__device__ __forceinline__ Field<P> team_shfl(unsigned mask, const Field<P>& v, int src) {
  Field<P> r;
#pragma unroll
  for (int i = 0; i < Field<P>::L; i++) r.v[i] = __shfl_sync(mask, v.v[i], src);
  return r;
}
template <class P>
__device__ __forceinline__ Field2<P> team_shfl(unsigned mask, const Field2<P>& v, int src) {
  return Field2<P>{team_shfl(mask, v.a, src), team_shfl(mask, v.b, src)};
}

// Lanes that share one group operation: 4 — one product of the formula step per lane (roles 0..3) — or, for G2, 16 = 4
// roles x 4 parts, where a role's Fp2 product is split into its three Karatsuba Fp products (a a', b b',
// (a + b)(a' + b'); the fourth part repeats the third), so that a product round is ONE Fp product deep instead of
// three.  16-lane teams finish an addition in ~8 us against ~18 us, but a warp holds two of them instead of eight:
// they win on levels of up to ~2400 items and lose above (profiles/r2_g2_teams.md), so the width is a template
// parameter of the operations and msm_tail picks it per level.  TeamWidth<F> is the default for the narrow steps.
template <class F> struct TeamWidth;
template <class P> struct TeamWidth<Field<P>> { static constexpr int LANES = 4, SHIFT = 2; };
template <class P> struct TeamWidth<Field2<P>> { static constexpr int LANES = 16, SHIFT = 4; };

template <class F, int LANES_ = TeamWidth<F>::LANES>
struct Team {
  static_assert(LANES_ == 4 || LANES_ == 16, "teams are 4 or 16 lanes wide");
  static constexpr int LANES = LANES_, SHIFT = LANES_ == 4 ? 2 : 4, PER_WARP = 32 / LANES_;
  int lane, role, part, base;
  unsigned mask;
  __device__ __forceinline__ Team() {
    lane = threadIdx.x & 31;
    base = lane & ~(LANES - 1);
    mask = ((1u << LANES) - 1u) << base;
    role = LANES == 4 ? (lane & 3) : ((lane >> 2) & 3);
    part = LANES == 4 ? 0 : (lane & 3);
  }
  __device__ __forceinline__ bool lead() const { return (lane & (LANES - 1)) == 0; }
};

// One product per ROLE: every lane passes the operands of its role and gets back its PIECE of that product ...
template <class F>
__device__ __forceinline__ F team_piece(const Team<F, 4>&, const F& x, const F& y) {
  return F::mul(x, y);  // the whole product (for Fp2: three dependent Fp products in this lane)
}
// 16 lanes: part 0 / 1 / 2 of a role multiply a a' / b b' / (a + b)(a' + b'); the Karatsuba recombination is done ONCE
// here, by the role's own lanes (two shuffles and two subtractions for all four roles at the same time), so that
// the piece of part 0 is the real and the piece of part 1 the imaginary component of the role's product — a consumer
// then fetches a product with two shuffles and no arithmetic (recombining at every consumer cost ~0.5 us per round).
template <class P>
__device__ __forceinline__ Field<P> team_piece(const Team<Field2<P>, 16>& tm, const Field2<P>& x, const Field2<P>& y) {
  using Base = Field<P>;
  const Base u = tm.part == 0 ? x.a : tm.part == 1 ? x.b : Base::add(x.a, x.b);
  const Base v = tm.part == 0 ? y.a : tm.part == 1 ? y.b : Base::add(y.a, y.b);
  const Base m = Base::mul(u, v);
  const Base other = team_shfl(tm.mask, m, tm.lane ^ 1);         // part 0 gets b b', part 1 gets a a'
  const Base third = team_shfl(tm.mask, m, (tm.lane & ~3) | 2);  // (a + b)(a' + b')
  // part 0:  a a' - b b' - 0 ;  part 1:  (a + b)(a' + b') - a a' - b b'   (parts 2 and 3: unused)
  return Base::sub(Base::sub(tm.part == 0 ? m : third, other), tm.part == 0 ? Base::zero() : m);
}
// ... and the complete product of role r, on every lane of the team
template <class F>
__device__ __forceinline__ F team_product(const Team<F, 4>& tm, const F& piece, int r) {
  return team_shfl(tm.mask, piece, tm.base + r);
}
template <class P>
__device__ __forceinline__ Field2<P> team_product(const Team<Field2<P>, 16>& tm, const Field<P>& piece, int r) {
  return Field2<P>{team_shfl(tm.mask, piece, tm.base + 4 * r), team_shfl(tm.mask, piece, tm.base + 4 * r + 1)};
}

// a + b (add-2008-s, same case analysis as XYZZ::add); a, b and the result are replicated on the team
template <class F, int LANES = TeamWidth<F>::LANES>
__device__ __noinline__ XYZZ<F> team_add(const XYZZ<F>& a, const XYZZ<F>& b) {
  const Team<F, LANES> tm;
  const int role = tm.role;
  if (b.is_inf()) return a;
  if (a.is_inf()) return b;
  // round 1:  U1 = X1 ZZ2 | U2 = X2 ZZ1 | S1 = Y1 ZZZ2 | S2 = Y2 ZZZ1
  const auto p1 = team_piece(tm, role == 0 ? a.X : role == 1 ? b.X : role == 2 ? a.Y : b.Y,
                             role == 0 ? b.ZZ : role == 1 ? a.ZZ : role == 2 ? b.ZZZ : a.ZZZ);
  // round 2:  ZZ1 ZZ2 (even roles) | ZZZ1 ZZZ2 (odd roles)
  const auto p2 = team_piece(tm, (role & 1) ? a.ZZZ : a.ZZ, (role & 1) ? b.ZZZ : b.ZZ);
  const F U1 = team_product(tm, p1, 0), S1 = team_product(tm, p1, 2);
  const F Pd = F::sub(team_product(tm, p1, 1), U1), R = F::sub(team_product(tm, p1, 3), S1);
  if (Pd.is_zero()) {
    XYZZ<F> s = a;
    if (R.is_zero()) s.dbl();
    else s = XYZZ<F>::inf();
    return s;
  }
  // round 3:  PP (even roles) | RR (odd roles)
  const auto p3 = team_piece(tm, (role & 1) ? R : Pd, (role & 1) ? R : Pd);
  const F PP = team_product(tm, p3, 0), RR = team_product(tm, p3, 1);
  // round 4:  PPP = P PP | Q = U1 PP | ZZ3 = ZZ1 ZZ2 PP | (role 3 repeats role 2)
  const F ZZ12 = team_product(tm, p2, 0), ZZZ12 = team_product(tm, p2, 1);
  const auto p4 = team_piece(tm, role == 0 ? Pd : role == 1 ? U1 : ZZ12, PP);
  const F PPP = team_product(tm, p4, 0), Q = team_product(tm, p4, 1);
  XYZZ<F> o;
  o.ZZ = team_product(tm, p4, 2);
  o.X = F::sub(F::sub(RR, PPP), F::dbl(Q));
  // round 5:  R (Q - X3) | S1 PPP | ZZZ1 ZZZ2 PPP
  const auto p5 = team_piece(tm, role == 0 ? R : role == 1 ? S1 : ZZZ12, role == 0 ? F::sub(Q, o.X) : PPP);
  o.Y = F::sub(team_product(tm, p5, 0), team_product(tm, p5, 1));
  o.ZZZ = team_product(tm, p5, 2);
  return o;
}

// 2 a (dbl-2008-s-1 with a = 0, as XYZZ::dbl)
template <class F, int LANES = TeamWidth<F>::LANES>
__device__ __noinline__ XYZZ<F> team_dbl(const XYZZ<F>& a) {
  const Team<F, LANES> tm;
  const int role = tm.role;
  if (a.is_inf()) return a;
  const F U = F::dbl(a.Y);
  // round 1:  V = U^2 (even roles) | X^2 (odd roles)
  const auto p1 = team_piece(tm, (role & 1) ? a.X : U, (role & 1) ? a.X : U);
  const F V = team_product(tm, p1, 0), xx = team_product(tm, p1, 1);
  const F M = F::add(F::dbl(xx), xx);
  // round 2:  W = U V | S = X V | ZZ3 = ZZ V | M^2
  const auto p2 = team_piece(tm, role == 0 ? U : role == 1 ? a.X : role == 2 ? a.ZZ : M, role == 3 ? M : V);
  const F W = team_product(tm, p2, 0), S = team_product(tm, p2, 1), MM = team_product(tm, p2, 3);
  XYZZ<F> o;
  o.ZZ = team_product(tm, p2, 2);
  o.X = F::sub(MM, F::dbl(S));
  // round 3:  M (S - X3) | W Y | W ZZZ
  const auto p3 = team_piece(tm, role == 0 ? M : W, role == 0 ? F::sub(S, o.X) : role == 1 ? a.Y : a.ZZZ);
  o.Y = F::sub(team_product(tm, p3, 0), team_product(tm, p3, 1));
  o.ZZZ = team_product(tm, p3, 2);
  return o;
}

template <class F, int LANES = TeamWidth<F>::LANES>
__device__ XYZZ<F> team_small_mul(const XYZZ<F>& p, uint32_t k) {
  XYZZ<F> m = XYZZ<F>::inf();
  if (k == 0 || p.is_inf()) return m;
  for (int bit = top_bit_u32(k); bit >= 0; bit--) {
    m = team_dbl<F, LANES>(m);
    if ((k >> bit) & 1) m = team_add<F, LANES>(m, p);
  }
  return m;
}

// Team counterparts of the msm.cuh reduction kernels: item = global thread / lanes.  The separate launches serve the
// WIDE levels (msm_tail takes the narrow ones), where lane-microseconds per operation count, not the length of one
// operation: they run on WIDE_LANES = 4 lanes for both groups.  msm_finish_team and sum_points_team are narrow.
constexpr int WIDE_LANES = 4, WIDE_SHIFT = 2;
#define KZG_TEAM_LOOP_W(var, total, SHIFT)                                                          \
  for (uint64_t var = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> (SHIFT); var < (total);  \
       var += ((uint64_t)gridDim.x * blockDim.x) >> (SHIFT))
#define KZG_TEAM_LOOP(var, total) KZG_TEAM_LOOP_W(var, total, WIDE_SHIFT)

template <class F>
__global__ void __launch_bounds__(128)
msm_fold_level_team(int level, const uint32_t* __restrict__ ntasks_ptr, const uint32_t* __restrict__ max_nt,
                    const MsmSlotSeg* __restrict__ slot_seg, XYZZ<F>* __restrict__ partial) {
  const uint32_t stride = 1u << level;
  if (stride >= *max_nt) return;
  const uint32_t ntasks = *ntasks_ptr;
  KZG_TEAM_LOOP(q, ntasks) {
    const MsmSlotSeg sg = slot_seg[q];
    if ((sg.k & (2u * stride - 1u)) == 0u && sg.k + stride < sg.nt) {
      const XYZZ<F> a = partial[q], b = partial[q + stride];
      const XYZZ<F> s = team_add<F, WIDE_LANES>(a, b);
      __syncwarp(Team<F, WIDE_LANES>().mask);  // every lane has read partial[q] before lane 0 overwrites it
      if (Team<F, WIDE_LANES>().lead()) partial[q] = s;
    }
  }
}

template <class F>
__global__ void __launch_bounds__(128)
msm_bucket_pairs_team(uint64_t nb, const uint32_t* __restrict__ task_first, const XYZZ<F>* __restrict__ partial,
                      BucketJob<F> j0, BucketJob<F> j1) {
  const BucketJob<F>& jb = blockIdx.y ? j1 : j0;
  KZG_TEAM_LOOP(t, jb.total) {
    const uint64_t P = jb.len >= 2 ? jb.len / 2 : 1;
    const uint64_t p = t % P, gs = t / P;
    const uint64_t s = gs % jb.nsums, g = gs / jb.nsums;
    const uint64_t b0 = g * nb + s * jb.sstride + 2 * p * jb.jstride;
    XYZZ<F> acc = bucket_value<F>(task_first, partial, b0);
    if (jb.len >= 2) acc = team_add<F, WIDE_LANES>(acc, bucket_value<F>(task_first, partial, b0 + jb.jstride));
    if (Team<F, WIDE_LANES>().lead()) jb.out[t] = acc;
  }
}

template <class F>
__global__ void __launch_bounds__(128, 4) msm_pair_sum_team(PairJob<F> j0, PairJob<F> j1) {
  const PairJob<F>& jb = blockIdx.y ? j1 : j0;
  KZG_TEAM_LOOP(i, jb.n_out) {
    const XYZZ<F> a = jb.in[2 * i], b = jb.in[2 * i + 1];
    const XYZZ<F> s = team_add<F, WIDE_LANES>(a, b);
    if (Team<F, WIDE_LANES>().lead()) jb.out[i] = s;
  }
}

template <class F>
__global__ void __launch_bounds__(128) msm_weighted_chunks_team(WeightedJob<F> j0, WeightedJob<F> j1) {
  const WeightedJob<F>& jb = blockIdx.y ? j1 : j0;
  KZG_TEAM_LOOP(t, jb.total) {  // same sums as msm_weighted_chunk_thread
    const uint64_t m = jb.m;
    const uint64_t P = m >= WCHUNK ? m / WCHUNK : 1;
    const uint64_t p = t % P, g = t / P;
    const uint64_t j0i = p * WCHUNK, j1i = (j0i + WCHUNK < m) ? j0i + WCHUNK : m;
    XYZZ<F> run = XYZZ<F>::inf(), acc = XYZZ<F>::inf();
    for (uint64_t j = j1i; j > j0i; j--) {
      const XYZZ<F> v = jb.in[g * m + j - 1];
      run = team_add<F, WIDE_LANES>(run, v);
      acc = team_add<F, WIDE_LANES>(acc, run);
    }
    const uint32_t k = (uint32_t)j0i + jb.woff;
    if (k >= 1) acc = team_add<F, WIDE_LANES>(acc, team_small_mul<F, WIDE_LANES>(run, k - 1));
    else {
      run.negate();
      acc = team_add<F, WIDE_LANES>(acc, run);
    }
    if (Team<F, WIDE_LANES>().lead()) jb.out[t] = acc;
  }
}

template <class F>
__global__ void __launch_bounds__(128)
msm_group_combine_team(uint64_t groups, int h, const XYZZ<F>* __restrict__ t1, const XYZZ<F>* __restrict__ t2,
                       XYZZ<F>* __restrict__ group_out) {
  KZG_TEAM_LOOP(g, groups) {
    XYZZ<F> acc = t1[g];
    for (int k = 0; k < h; k++) acc = team_dbl<F, WIDE_LANES>(acc);
    const XYZZ<F> b = t2[g];
    acc = team_add<F, WIDE_LANES>(acc, b);
    if (Team<F, WIDE_LANES>().lead()) group_out[g] = acc;
  }
}

// msm_finish with the Horner chain over the windows (no table: W - 1 rounds of c doublings + one addition, all
// dependent) on a team; lane 0 then normalises and writes the bytes as msm_finish_thread does.
template <class F>
__global__ void __launch_bounds__(128)
msm_finish_team(uint64_t batch, int groups_per_msm, int c, const XYZZ<F>* __restrict__ group_out,
                uint8_t* __restrict__ out_bytes, int* __restrict__ is_inf, int modbytes, XYZZ<F>* __restrict__ out_xyzz,
                int want_affine) {
  KZG_TEAM_LOOP_W(m, batch, TeamWidth<F>::SHIFT) {
    XYZZ<F> acc = group_out[m * groups_per_msm + (groups_per_msm - 1)];
    for (int g = groups_per_msm - 2; g >= 0; g--) {
      for (int k = 0; k < c; k++) acc = team_dbl<F>(acc);
      const XYZZ<F> v = group_out[m * groups_per_msm + g];
      acc = team_add<F>(acc, v);
    }
    if (Team<F>().lead()) {
      if (out_xyzz) out_xyzz[m] = acc;
      const int nbytes = CoordCount<F>::value * modbytes;
      if (!want_affine) {
        is_inf[m] = acc.is_inf() ? 1 : 0;
      } else if (acc.is_inf()) {
        for (int j = 0; j < nbytes; j++) out_bytes[m * nbytes + j] = 0;
        is_inf[m] = 1;
      } else {
        affine_to_bytes(acc.to_affine(), out_bytes + m * nbytes, modbytes);
        is_inf[m] = 0;
      }
    }
  }
}

// sum of n projective points (the partial sums of a point-range split gathered from all ranks), then affine bytes:
// 128 teams take strided slices, a shared-memory tree joins them (log2 levels of one team addition each).
template <class F>
__global__ void __launch_bounds__(512)
sum_points_team(uint64_t n, const XYZZ<F>* __restrict__ pts, uint8_t* __restrict__ out_bytes, int* __restrict__ is_inf,
                int modbytes) {
  constexpr int NT = 512 >> TeamWidth<F>::SHIFT;  // teams in the CTA: 128 of 4 lanes (G1)
  __shared__ XYZZ<F> sh[NT];
  const int team = threadIdx.x >> TeamWidth<F>::SHIFT;
  const unsigned mask = Team<F>().mask;
  XYZZ<F> acc = XYZZ<F>::inf();
  for (uint64_t i = team; i < n; i += NT) {
    const XYZZ<F> v = pts[i];
    acc = team_add<F>(acc, v);
  }
  if (Team<F>().lead()) sh[team] = acc;
  __syncthreads();
  for (int stride = NT / 2; stride >= 1; stride >>= 1) {
    if (team < stride) {
      const XYZZ<F> a = sh[team], b = sh[team + stride];
      const XYZZ<F> s = team_add<F>(a, b);
      __syncwarp(mask);
      if (Team<F>().lead()) sh[team] = s;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const XYZZ<F> total = sh[0];
    const int nbytes = CoordCount<F>::value * modbytes;
    if (total.is_inf()) {
      for (int j = 0; j < nbytes; j++) out_bytes[j] = 0;
      is_inf[0] = 1;
    } else {
      affine_to_bytes(total.to_affine(), out_bytes, modbytes);
      is_inf[0] = 0;
    }
  }
}

}  // namespace kzgb200
#endif
