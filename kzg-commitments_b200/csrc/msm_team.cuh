// The latency-bound tail of the bucket reduction on 4-lane teams.
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

struct Team {
  int role, base;
  unsigned mask;
  __device__ __forceinline__ Team() {
    const int lane = threadIdx.x & 31;
    role = lane & 3;
    base = lane & ~3;
    mask = 0xFu << base;
  }
};

// a + b (add-2008-s, same case analysis as XYZZ::add); a, b and the result are replicated on the team
template <class F>
__device__ __noinline__ XYZZ<F> team_add(const XYZZ<F>& a, const XYZZ<F>& b) {
  const Team tm;
  const int role = tm.role, base = tm.base;
  const unsigned mask = tm.mask;
  if (b.is_inf()) return a;
  if (a.is_inf()) return b;
  // round 1:  U1 = X1 ZZ2 | U2 = X2 ZZ1 | S1 = Y1 ZZZ2 | S2 = Y2 ZZZ1
  const F m1 = F::mul(role == 0 ? a.X : role == 1 ? b.X : role == 2 ? a.Y : b.Y,
                      role == 0 ? b.ZZ : role == 1 ? a.ZZ : role == 2 ? b.ZZZ : a.ZZZ);
  // round 2:  ZZ1 ZZ2 (even lanes) | ZZZ1 ZZZ2 (odd lanes)
  const F m2 = F::mul((role & 1) ? a.ZZZ : a.ZZ, (role & 1) ? b.ZZZ : b.ZZ);
  // lanes 0,1: P = U2 - U1;  lanes 2,3: R = S2 - S1
  const F other = team_shfl(mask, m1, (threadIdx.x & 31) ^ 1);
  const F d = (role & 1) ? F::sub(m1, other) : F::sub(other, m1);
  const F Pd = team_shfl(mask, d, base), R = team_shfl(mask, d, base + 2);
  if (Pd.is_zero()) {
    XYZZ<F> s = a;
    if (R.is_zero()) s.dbl();
    else s = XYZZ<F>::inf();
    return s;
  }
  // round 3:  PP (lanes 0,1) | RR (lanes 2,3)
  const F sq = F::sqr(d);
  const F PP = team_shfl(mask, sq, base), RR = team_shfl(mask, sq, base + 2);
  // round 4:  PPP = P PP | Q = U1 PP | ZZ3 = ZZ1 ZZ2 PP | (lane 3 repeats lane 2)
  const F U1 = team_shfl(mask, m1, base), S1 = team_shfl(mask, m1, base + 2);
  const F ZZ12 = team_shfl(mask, m2, base), ZZZ12 = team_shfl(mask, m2, base + 1);
  const F m4 = F::mul(role == 0 ? Pd : role == 1 ? U1 : ZZ12, PP);
  const F PPP = team_shfl(mask, m4, base), Q = team_shfl(mask, m4, base + 1);
  XYZZ<F> o;
  o.ZZ = team_shfl(mask, m4, base + 2);
  o.X = F::sub(F::sub(RR, PPP), F::dbl(Q));
  // round 5:  R (Q - X3) | S1 PPP | ZZZ1 ZZZ2 PPP
  const F m5 = F::mul(role == 0 ? R : role == 1 ? S1 : ZZZ12, role == 0 ? F::sub(Q, o.X) : PPP);
  o.Y = F::sub(team_shfl(mask, m5, base), team_shfl(mask, m5, base + 1));
  o.ZZZ = team_shfl(mask, m5, base + 2);
  return o;
}

// 2 a (dbl-2008-s-1 with a = 0, as XYZZ::dbl)
template <class F>
__device__ __noinline__ XYZZ<F> team_dbl(const XYZZ<F>& a) {
  const Team tm;
  const int role = tm.role, base = tm.base;
  const unsigned mask = tm.mask;
  if (a.is_inf()) return a;
  const F U = F::dbl(a.Y);
  // round 1:  V = U^2 (even lanes) | X^2 (odd lanes)
  const F m1 = F::sqr((role & 1) ? a.X : U);
  const F V = team_shfl(mask, m1, base), xx = team_shfl(mask, m1, base + 1);
  const F M = F::add(F::dbl(xx), xx);
  // round 2:  W = U V | S = X V | ZZ3 = ZZ V | M^2
  const F m2 = F::mul(role == 0 ? U : role == 1 ? a.X : role == 2 ? a.ZZ : M, role == 3 ? M : V);
  const F W = team_shfl(mask, m2, base), S = team_shfl(mask, m2, base + 1), MM = team_shfl(mask, m2, base + 3);
  XYZZ<F> o;
  o.ZZ = team_shfl(mask, m2, base + 2);
  o.X = F::sub(MM, F::dbl(S));
  // round 3:  M (S - X3) | W Y | W ZZZ
  const F m3 = F::mul(role == 0 ? M : W, role == 0 ? F::sub(S, o.X) : role == 1 ? a.Y : a.ZZZ);
  o.Y = F::sub(team_shfl(mask, m3, base), team_shfl(mask, m3, base + 1));
  o.ZZZ = team_shfl(mask, m3, base + 2);
  return o;
}

template <class F>
__device__ XYZZ<F> team_small_mul(const XYZZ<F>& p, uint32_t k) {
  XYZZ<F> m = XYZZ<F>::inf();
  if (k == 0 || p.is_inf()) return m;
  for (int bit = top_bit_u32(k); bit >= 0; bit--) {
    m = team_dbl<F>(m);
    if ((k >> bit) & 1) m = team_add<F>(m, p);
  }
  return m;
}

// Team counterparts of the msm.cuh reduction kernels: item = global thread / 4.
#define KZG_TEAM_LOOP(var, total)                                                           \
  for (uint64_t var = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 2; var < (total); \
       var += ((uint64_t)gridDim.x * blockDim.x) >> 2)

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
      const XYZZ<F> s = team_add<F>(a, b);
      __syncwarp(Team().mask);  // every lane has read partial[q] before lane 0 overwrites it
      if ((threadIdx.x & 3) == 0) partial[q] = s;
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
    if (jb.len >= 2) acc = team_add<F>(acc, bucket_value<F>(task_first, partial, b0 + jb.jstride));
    if ((threadIdx.x & 3) == 0) jb.out[t] = acc;
  }
}

template <class F>
__global__ void __launch_bounds__(128, 4) msm_pair_sum_team(PairJob<F> j0, PairJob<F> j1) {
  const PairJob<F>& jb = blockIdx.y ? j1 : j0;
  KZG_TEAM_LOOP(i, jb.n_out) {
    const XYZZ<F> a = jb.in[2 * i], b = jb.in[2 * i + 1];
    const XYZZ<F> s = team_add<F>(a, b);
    if ((threadIdx.x & 3) == 0) jb.out[i] = s;
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
      run = team_add<F>(run, v);
      acc = team_add<F>(acc, run);
    }
    const uint32_t k = (uint32_t)j0i + jb.woff;
    if (k >= 1) acc = team_add<F>(acc, team_small_mul<F>(run, k - 1));
    else {
      run.negate();
      acc = team_add<F>(acc, run);
    }
    if ((threadIdx.x & 3) == 0) jb.out[t] = acc;
  }
}

template <class F>
__global__ void __launch_bounds__(128)
msm_group_combine_team(uint64_t groups, int h, const XYZZ<F>* __restrict__ t1, const XYZZ<F>* __restrict__ t2,
                       XYZZ<F>* __restrict__ group_out) {
  KZG_TEAM_LOOP(g, groups) {
    XYZZ<F> acc = t1[g];
    for (int k = 0; k < h; k++) acc = team_dbl<F>(acc);
    const XYZZ<F> b = t2[g];
    acc = team_add<F>(acc, b);
    if ((threadIdx.x & 3) == 0) group_out[g] = acc;
  }
}

// msm_finish with the Horner chain over the windows (no table: W - 1 rounds of c doublings + one addition, all
// dependent) on a team; lane 0 then normalises and writes the bytes as msm_finish_thread does.
template <class F>
__global__ void __launch_bounds__(128)
msm_finish_team(uint64_t batch, int groups_per_msm, int c, const XYZZ<F>* __restrict__ group_out,
                uint8_t* __restrict__ out_bytes, int* __restrict__ is_inf, int modbytes, XYZZ<F>* __restrict__ out_xyzz,
                int want_affine) {
  KZG_TEAM_LOOP(m, batch) {
    XYZZ<F> acc = group_out[m * groups_per_msm + (groups_per_msm - 1)];
    for (int g = groups_per_msm - 2; g >= 0; g--) {
      for (int k = 0; k < c; k++) acc = team_dbl<F>(acc);
      const XYZZ<F> v = group_out[m * groups_per_msm + g];
      acc = team_add<F>(acc, v);
    }
    if ((threadIdx.x & 3) == 0) {
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
  __shared__ XYZZ<F> sh[128];
  const int team = threadIdx.x >> 2;
  const unsigned mask = Team().mask;
  XYZZ<F> acc = XYZZ<F>::inf();
  for (uint64_t i = team; i < n; i += 128) {
    const XYZZ<F> v = pts[i];
    acc = team_add<F>(acc, v);
  }
  if ((threadIdx.x & 3) == 0) sh[team] = acc;
  __syncthreads();
  for (int stride = 64; stride >= 1; stride >>= 1) {
    if (team < stride) {
      const XYZZ<F> a = sh[team], b = sh[team + stride];
      const XYZZ<F> s = team_add<F>(a, b);
      __syncwarp(mask);
      if ((threadIdx.x & 3) == 0) sh[team] = s;
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
