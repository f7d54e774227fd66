// Persistent (cooperative) kernels for the latency-bound ends of the Pippenger pipeline.
//
// A small or medium MSM (one create_commit of degree 4096, the per-GPU share of a point-range split) spends most
// of its time in steps that cannot fill the machine: a dozen sort/scan launches of a few microseconds each in
// front of the bucket accumulation and ~35 reduction launches behind it (round 1: 2^12 points = 0.60 ms of which
// 0.05 ms accumulate; profiles/r1k_curves_and_sizes.md).  Here each end is ONE kernel with exactly one CTA per SM
// (cudaLaunchCooperativeKernel: all resident) that walks through the same steps separated by grid barriers:
//
//   msm_sort_small   zero -> digits + histogram -> bucket offsets (3-phase scan) -> scatter -> tasks per bucket
//                    (scan) -> task-length histogram -> its scan -> task placement            (was 13 launches)
//   msm_tail         fold of task partials -> 2-D bucket reduction (row / column trees, weighted chunks, trees)
//                    -> group combine -> Horner over windows, affine normalisation, bytes     (was up to ~40 launches)
//
// What bounds these steps (measured, profiles/r2a_*): ONE warp's Montgomery product already saturates its SM
// sub-partition's multiplier pipe (an IMAD.WIDE issues every 4 cycles per SMSP), so a dependent chain of group
// operations runs at full speed only with one warp per sub-partition; a second warp on the same sub-partition
// halves both.  The first version of this kernel packed the items of a narrow level into the first few CTAs
// (4 warps per sub-partition) and was SLOWER than the launches it replaced.  Hence the layout here:
//   * 148 CTAs x 128 threads, one CTA per SM (a large dynamic shared-memory request keeps a second one away),
//     i.e. exactly one warp per sub-partition;
//   * items are dealt out warp-major ACROSS the CTAs: a level of <= 1184 items keeps one warp per SM busy,
//     <= 4736 items one warp per sub-partition (4-lane teams, msm_team.cuh: 8 group operations per warp);
//   * levels wider than two such rounds run one group operation per THREAD instead (14 dependent products, but
//     32 per warp), which is the machine-filling form; still wider levels stay separate launches (ctx.cuh).
// msm_tail takes a [first_stage, last_stage] range so that ctx.cuh can place the wide launches between a
// fold-only call and the call that finishes the narrow levels.  Data passes between steps through global memory.
// The grid barrier is a generation counter in device memory: one atomic per CTA to arrive, the last arrival bumps
// the generation, the others poll it with ld.acquire and a short sleep (148 pollers hammering one L2 line was the
// other thing that made the first version slow); __threadfence on both sides gives the usual grid.sync() ordering
// and invalidates the SM's L1 before data written by other CTAs is read.
#pragma once
#include "grid_barrier.cuh"
#include "msm_team.cuh"

#if defined(__CUDACC__)

namespace kzgb200 {

constexpr int TAIL_THREADS = 128;          // one warp per SM sub-partition
constexpr int TAIL_SMEM_BYTES = 120 << 10; // dynamic shared memory requested only to keep the CTA alone on its SM

// Item distribution: virtual warp v = warp-in-CTA * gridDim + CTA, so consecutive groups of items land on
// different SMs first and on a second warp of the same SM only after every SM has one.
template <class F>
struct TailLane {
  uint64_t vwarp, nvwarps;
  int lane;
  __device__ __forceinline__ TailLane() {
    lane = threadIdx.x & 31;
    vwarp = (uint64_t)(threadIdx.x >> 5) * gridDim.x + blockIdx.x;
    nvwarps = (uint64_t)(blockDim.x >> 5) * gridDim.x;
  }
  template <int LANES> __device__ __forceinline__ uint64_t team() const { return vwarp * (32 / LANES) + (lane / LANES); }
  template <int LANES> __device__ __forceinline__ uint64_t nteams() const { return nvwarps * (32 / LANES); }
  __device__ __forceinline__ uint64_t thread() const { return vwarp * 32 + lane; }
  __device__ __forceinline__ uint64_t nthreads() const { return nvwarps * 32; }
  // How a level of `items` group operations runs: on 16-lane teams (G2 only: ~8 us per round, 1184 per round), on
  // 4-lane teams (G1 6.7 us / G2 ~18 us per round, 4736 per round) while that beats ONE round of one operation per
  // thread (G1 14.8 us / G2 ~45 us, 18944 per round), else one operation per thread (returns 0).
  __device__ __forceinline__ int mode(uint64_t items) const {
    if (TeamWidth<F>::LANES == 16 && items <= 2 * nteams<16>()) return 16;
    if (items <= 2 * nteams<4>()) return 4;
    return 0;
  }
};

enum TailStage : int {
  TAIL_FOLD = 0, TAIL_BUCKET_PAIRS = 1, TAIL_TREES1 = 2, TAIL_WEIGHTED = 3, TAIL_TREES2 = 4, TAIL_COMBINE = 5, TAIL_FINISH = 6
};

template <class F>
struct TailArgs {
  int first_stage, last_stage;
  GridBarrier* barrier;
  // fold
  const uint32_t* ntasks_ptr;
  const uint32_t* max_nt;
  const MsmSlotSeg* slot_seg;
  XYZZ<F>* partial;
  // bucket values
  const uint32_t* task_first;
  uint64_t nb;                 // buckets per group
  // 2-D reduction: ping-pong buffers of the row sums (rin/rout) and column sums (cin/cout)
  XYZZ<F>*rin, *rout, *cin, *cout;
  uint64_t groups, R, Cn;      // bucket groups, rows and columns per group
  uint64_t r_len, c_len;       // values still to add per row / column sum when the kernel enters at TAIL_TREES1
                               // (or per group at TAIL_TREES2)
  int h;                       // Cn = 2^h
  XYZZ<F>* group_out;
  // finish
  uint64_t batch;
  int groups_per_msm, c, modbytes, want_affine;
  uint8_t* out_bytes;
  int* is_inf;
  XYZZ<F>* out_xyzz;
};

// ---- the steps on teams of LANES lanes ---------------------------------------------------------------------
// one level of two pairwise trees side by side: out[i] = in[2i] + in[2i+1]
template <class F, int LANES>
__device__ __forceinline__ void tail_pair_level_teams(const TailLane<F>& tl, const XYZZ<F>* in0, XYZZ<F>* out0, uint64_t n0,
                                                      const XYZZ<F>* in1, XYZZ<F>* out1, uint64_t n1) {
  const bool lead = Team<F, LANES>().lead();
  for (uint64_t t = tl.template team<LANES>(); t < n0 + n1; t += tl.template nteams<LANES>()) {
    const bool second = t >= n0;
    const uint64_t i = second ? t - n0 : t;
    const XYZZ<F>* in = second ? in1 : in0;
    const XYZZ<F> a = in[2 * i], b = in[2 * i + 1];
    const XYZZ<F> s = team_add<F, LANES>(a, b);
    if (lead) (second ? out1 : out0)[i] = s;
  }
}
template <class F>
__device__ __forceinline__ void tail_pair_level(const TailLane<F>& tl, const XYZZ<F>* in0, XYZZ<F>* out0, uint64_t n0,
                                                const XYZZ<F>* in1, XYZZ<F>* out1, uint64_t n1) {
  const int mode = tl.mode(n0 + n1);
  if (mode == 4) {
    tail_pair_level_teams<F, 4>(tl, in0, out0, n0, in1, out1, n1);
  } else if (mode == 16) {
    if constexpr (TeamWidth<F>::LANES == 16) tail_pair_level_teams<F, 16>(tl, in0, out0, n0, in1, out1, n1);
  } else {
    for (uint64_t t = tl.thread(); t < n0 + n1; t += tl.nthreads()) {
      const bool second = t >= n0;
      const uint64_t i = second ? t - n0 : t;
      const XYZZ<F>* in = second ? in1 : in0;
      XYZZ<F> acc = in[2 * i];
      acc.add(in[2 * i + 1]);
      (second ? out1 : out0)[i] = acc;
    }
  }
}

// halve both trees until one value per sum is left; a grid barrier after every level
template <class F>
__device__ __forceinline__ void tail_trees(const TailLane<F>& tl, GridBarrier* bar, XYZZ<F>*& in0, XYZZ<F>*& out0, uint64_t sums0,
                                           uint64_t& len0, XYZZ<F>*& in1, XYZZ<F>*& out1, uint64_t sums1, uint64_t& len1) {
  while (len0 > 1 || len1 > 1) {
    tail_pair_level<F>(tl, in0, out0, len0 > 1 ? sums0 * (len0 / 2) : 0, in1, out1, len1 > 1 ? sums1 * (len1 / 2) : 0);
    if (len0 > 1) { XYZZ<F>* t = in0; in0 = out0; out0 = t; len0 /= 2; }
    if (len1 > 1) { XYZZ<F>* t = in1; in1 = out1; out1 = t; len1 /= 2; }
    grid_sync(bar);
  }
}

// One level of the segmented fold over the task partials (msm_fold_level_team): a warp scans 32 consecutive slots at a
// time (coalesced) and its teams then take the few that have work, one per team at a time.
template <class F, int LANES>
__device__ __forceinline__ void tail_fold_level(const TailLane<F>& tl, const TailArgs<F>& a, uint32_t stride, uint32_t ntasks) {
  const Team<F, LANES> tm;
  constexpr int PER_WARP = 32 / LANES;
  for (uint64_t q0 = tl.vwarp * 32; q0 < ntasks; q0 += tl.nvwarps * 32) {
    const uint64_t q = q0 + tl.lane;
    bool work = false;
    if (q < ntasks) {
      const MsmSlotSeg sg = a.slot_seg[q];
      work = (sg.k & (2u * stride - 1u)) == 0u && sg.k + stride < sg.nt;
    }
    unsigned pending = __ballot_sync(0xffffffffu, work);
    while (pending) {
      const int my = tl.lane / LANES;
      unsigned m = pending;
      int pick = -1;
      for (int k = 0; k <= my && m; k++) {  // the (my+1)-th set bit
        pick = __ffs(m) - 1;
        m &= m - 1;
        if (k < my) pick = -1;
      }
      if (pick >= 0) {
        const uint64_t qq = q0 + pick;
        const XYZZ<F> x = a.partial[qq], y = a.partial[qq + stride];
        const XYZZ<F> s = team_add<F, LANES>(x, y);
        __syncwarp(tm.mask);  // every lane has read partial[qq] before the leader overwrites it
        if (tm.lead()) a.partial[qq] = s;
      }
      for (int k = 0; k < PER_WARP && pending; k++) pending &= pending - 1;  // drop the ones that were taken
    }
  }
}

// level 0 of the row and column trees, reading bucket values through task_first (msm_bucket_pairs)
template <class F, int LANES>
__device__ __forceinline__ void tail_bucket_pairs_teams(const TailLane<F>& tl, const TailArgs<F>& a, const BucketJob<F>& jr,
                                                        const BucketJob<F>& jc) {
  const bool lead = Team<F, LANES>().lead();
  const uint64_t total = jr.total + jc.total;
  for (uint64_t t0 = tl.template team<LANES>(); t0 < total; t0 += tl.template nteams<LANES>()) {
    const bool second = t0 >= jr.total;
    const BucketJob<F>& jb = second ? jc : jr;
    const uint64_t t = second ? t0 - jr.total : t0;
    const uint64_t P = jb.len >= 2 ? jb.len / 2 : 1;
    const uint64_t p = t % P, gs = t / P;
    const uint64_t s = gs % jb.nsums, g = gs / jb.nsums;
    const uint64_t b0 = g * a.nb + s * jb.sstride + 2 * p * jb.jstride;
    XYZZ<F> acc = bucket_value<F>(a.task_first, a.partial, b0);
    if (jb.len >= 2) acc = team_add<F, LANES>(acc, bucket_value<F>(a.task_first, a.partial, b0 + jb.jstride));
    if (lead) jb.out[t] = acc;
  }
}

// weighted chunks (msm_weighted_chunks): rows and columns side by side
template <class F, int LANES>
__device__ __forceinline__ void tail_weighted_teams(const TailLane<F>& tl, const WeightedJob<F>& wr, const WeightedJob<F>& wc) {
  const bool lead = Team<F, LANES>().lead();
  const uint64_t total = wr.total + wc.total;
  for (uint64_t t0 = tl.template team<LANES>(); t0 < total; t0 += tl.template nteams<LANES>()) {
    const bool second = t0 >= wr.total;
    const WeightedJob<F>& jb = second ? wc : wr;
    const uint64_t t = second ? t0 - wr.total : t0;
    const uint64_t m = jb.m;
    const uint64_t P = m >= WCHUNK ? m / WCHUNK : 1;
    const uint64_t p = t % P, g = t / P;
    const uint64_t j0i = p * WCHUNK, j1i = (j0i + WCHUNK < m) ? j0i + WCHUNK : m;
    XYZZ<F> run = XYZZ<F>::inf(), acc = XYZZ<F>::inf();
    for (uint64_t j = j1i; j > j0i; j--) {
      const XYZZ<F> v = jb.in[g * m + j - 1];
      run = team_add<F, LANES>(run, v);
      acc = team_add<F, LANES>(acc, run);
    }
    const uint32_t k = (uint32_t)j0i + jb.woff;
    if (k >= 1) acc = team_add<F, LANES>(acc, team_small_mul<F, LANES>(run, k - 1));
    else {
      run.negate();
      acc = team_add<F, LANES>(acc, run);
    }
    if (lead) jb.out[t] = acc;
  }
}

// group_out[g] = 2^h * T1[g] + T2[g]
template <class F, int LANES>
__device__ __forceinline__ void tail_combine_teams(const TailLane<F>& tl, const TailArgs<F>& a, const XYZZ<F>* rin, const XYZZ<F>* cin) {
  const bool lead = Team<F, LANES>().lead();
  for (uint64_t g = tl.template team<LANES>(); g < a.groups; g += tl.template nteams<LANES>()) {
    XYZZ<F> acc = rin[g];
    for (int k = 0; k < a.h; k++) acc = team_dbl<F, LANES>(acc);
    const XYZZ<F> b = cin[g];
    acc = team_add<F, LANES>(acc, b);
    if (lead) a.group_out[g] = acc;
  }
}

// dispatch on the lane count a level runs with; `thread_mode` is the one-operation-per-thread form
#define KZG_TAIL_DISPATCH(mode_expr, call4, call16, thread_mode)                \
  do {                                                                          \
    const int mode__ = (mode_expr);                                             \
    if (mode__ == 4) { call4; }                                                 \
    else if (mode__ == 16) { if constexpr (TeamWidth<F>::LANES == 16) { call16; } } \
    else { thread_mode; }                                                       \
  } while (0)

// ---- the steps on XYZZ values, every level in the form TailLane::mode picks (shared with msm_tail_proj.cuh, which
// ---- uses them for the levels that are too wide for its projective teams)
template <class F>
__device__ __forceinline__ void tail_fold_xyzz(const TailLane<F>& tl, const TailArgs<F>& a) {
  // Segmented binary tree over the task partials of each bucket, all levels in one launch.  At most ntasks / (2 stride)
  // pairs exist at a level; the fold has no one-operation-per-thread form, so wide levels use the 4-lane teams.
  const uint32_t max_nt = *a.max_nt, ntasks = *a.ntasks_ptr;
  for (uint32_t stride = 1; stride < max_nt; stride <<= 1) {
    const int mode = tl.mode(ntasks / (2 * stride) + 1);
    if (mode == 16) {
      if constexpr (TeamWidth<F>::LANES == 16) tail_fold_level<F, 16>(tl, a, stride, ntasks);
    } else {
      tail_fold_level<F, 4>(tl, a, stride, ntasks);
    }
    grid_sync(a.barrier);
  }
}
template <class F>
__device__ __forceinline__ void tail_bucket_pairs_xyzz(const TailLane<F>& tl, const TailArgs<F>& a, const BucketJob<F>& jr,
                                                       const BucketJob<F>& jc) {
  const uint64_t total = jr.total + jc.total;
  KZG_TAIL_DISPATCH(tl.mode(total), (tail_bucket_pairs_teams<F, 4>(tl, a, jr, jc)), (tail_bucket_pairs_teams<F, 16>(tl, a, jr, jc)),
                    for (uint64_t t0 = tl.thread(); t0 < total; t0 += tl.nthreads()) {
                      const bool second = t0 >= jr.total;
                      msm_bucket_pair_thread<F>(second ? jc : jr, a.nb, a.task_first, a.partial, second ? t0 - jr.total : t0);
                    });
}
template <class F>
__device__ __forceinline__ void tail_weighted_xyzz(const TailLane<F>& tl, const WeightedJob<F>& wr, const WeightedJob<F>& wc) {
  const uint64_t total = wr.total + wc.total;
  KZG_TAIL_DISPATCH(tl.mode(total), (tail_weighted_teams<F, 4>(tl, wr, wc)), (tail_weighted_teams<F, 16>(tl, wr, wc)),
                    for (uint64_t t0 = tl.thread(); t0 < total; t0 += tl.nthreads()) {
                      const bool second = t0 >= wr.total;
                      msm_weighted_chunk_thread<F>(second ? wc : wr, second ? t0 - wr.total : t0);
                    });
}
template <class F>
__device__ __forceinline__ void tail_combine_xyzz(const TailLane<F>& tl, const TailArgs<F>& a, const XYZZ<F>* rin, const XYZZ<F>* cin) {
  KZG_TAIL_DISPATCH(tl.mode(a.groups), (tail_combine_teams<F, 4>(tl, a, rin, cin)), (tail_combine_teams<F, 16>(tl, a, rin, cin)),
                    for (uint64_t g = tl.thread(); g < a.groups; g += tl.nthreads())
                      msm_group_combine_thread<F>(a.h, rin, cin, a.group_out, g););
}

template <class F>
__global__ void __launch_bounds__(TAIL_THREADS, 1) msm_tail(TailArgs<F> a) {
  const TailLane<F> tl;
  constexpr int NARROW = TeamWidth<F>::LANES;  // team width of the steps that are always narrow (finish)
  XYZZ<F>*rin = a.rin, *rout = a.rout, *cin = a.cin, *cout = a.cout;
  uint64_t r_len = a.r_len, c_len = a.c_len;

  if (a.first_stage <= TAIL_FOLD && a.last_stage >= TAIL_FOLD) tail_fold_xyzz<F>(tl, a);
  if (a.first_stage <= TAIL_BUCKET_PAIRS && a.last_stage >= TAIL_BUCKET_PAIRS) {
    const uint64_t Lr = a.Cn >= 2 ? a.Cn / 2 : 1, Lc = a.R >= 2 ? a.R / 2 : 1;
    const BucketJob<F> jr{rin, a.R, a.Cn, a.Cn, 1, a.groups * a.R * Lr};
    const BucketJob<F> jc{cin, a.Cn, a.R, 1, a.Cn, a.groups * a.Cn * Lc};
    tail_bucket_pairs_xyzz<F>(tl, a, jr, jc);
    r_len = Lr;
    c_len = Lc;
    grid_sync(a.barrier);
  }
  if (a.first_stage <= TAIL_TREES1 && a.last_stage >= TAIL_TREES1)
    tail_trees<F>(tl, a.barrier, rin, rout, a.groups * a.R, r_len, cin, cout, a.groups * a.Cn, c_len);
  if (a.first_stage <= TAIL_WEIGHTED && a.last_stage >= TAIL_WEIGHTED) {
    // rin = row sums [groups * R], cin = column sums [groups * Cn]
    const uint64_t Pr = a.R >= WCHUNK ? a.R / WCHUNK : 1, Pc = a.Cn >= WCHUNK ? a.Cn / WCHUNK : 1;
    const WeightedJob<F> wr{rin, rout, a.R, a.groups * Pr, 0u};
    const WeightedJob<F> wc{cin, cout, a.Cn, a.groups * Pc, 1u};
    tail_weighted_xyzz<F>(tl, wr, wc);
    { XYZZ<F>* t = rin; rin = rout; rout = t; }
    { XYZZ<F>* t = cin; cin = cout; cout = t; }
    r_len = Pr;
    c_len = Pc;
    grid_sync(a.barrier);
  }
  if (a.first_stage <= TAIL_TREES2 && a.last_stage >= TAIL_TREES2)
    tail_trees<F>(tl, a.barrier, rin, rout, a.groups, r_len, cin, cout, a.groups, c_len);
  if (a.first_stage <= TAIL_COMBINE && a.last_stage >= TAIL_COMBINE) {
    tail_combine_xyzz<F>(tl, a, rin, cin);
    if (a.last_stage >= TAIL_FINISH) grid_sync(a.barrier);
  }
  if (a.first_stage <= TAIL_FINISH && a.last_stage >= TAIL_FINISH) {
    // one MSM per WARP (its first team): the inversion behind to_affine is a data-dependent loop, and different
    // inversions in one warp would serialise
    if (tl.lane < NARROW) {
      const bool lead = tl.lane == 0;
      for (uint64_t m = tl.vwarp; m < a.batch; m += tl.nvwarps) {
        XYZZ<F> acc = a.group_out[m * a.groups_per_msm + (a.groups_per_msm - 1)];
        for (int g = a.groups_per_msm - 2; g >= 0; g--) {
          for (int k = 0; k < a.c; k++) acc = team_dbl<F, NARROW>(acc);
          const XYZZ<F> v = a.group_out[m * a.groups_per_msm + g];
          acc = team_add<F, NARROW>(acc, v);
        }
        if (lead) {
          if (a.out_xyzz) a.out_xyzz[m] = acc;
          const int nbytes = CoordCount<F>::value * a.modbytes;
          if (!a.want_affine) {
            a.is_inf[m] = acc.is_inf() ? 1 : 0;
          } else if (acc.is_inf()) {
            for (int j = 0; j < nbytes; j++) a.out_bytes[m * nbytes + j] = 0;
            a.is_inf[m] = 1;
          } else {
            affine_to_bytes(acc.to_affine(), a.out_bytes + m * nbytes, a.modbytes);
            a.is_inf[m] = 0;
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// The sort end: scalars -> signed digits -> records grouped by bucket -> tasks ordered by length, one launch.
struct SortArgs {
  MsmShape sh;
  GridBarrier* barrier;
  const uint8_t* scalars;
  int32_t* digits;
  uint32_t* ranks;
  uint32_t* counts;      // [nbuckets + 1]
  uint32_t* starts;      // [nbuckets + 1]
  uint32_t* task_first;  // [nbuckets + 1]
  uint32_t* sorted;
  uint32_t* tile_sums;   // [ntiles + 1]
  uint32_t* hist;        // task-length histogram block: hist | hist_start | cursor (ctx.cuh lays it out)
  uint32_t* hist_start;
  uint32_t* hist_cursor;
  MsmTask* tasks;
  MsmSlotSeg* slot_seg;
  uint64_t nbuckets;
  int task_chunk;        // records per task (the scan divides by it)
};

static __device__ __forceinline__ void sort_scan(GridBarrier* bar, uint64_t n, int chunk, const uint32_t* in, uint32_t* tile_sums,
                                                 uint32_t* out) {
  const uint64_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) scan_tile_sum_body(n, chunk, in, tile_sums, tile);
  grid_sync(bar);
  if (blockIdx.x == 0) scan_tiles_body(ntiles, tile_sums);
  grid_sync(bar);
  for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
    scan_apply_body(n, chunk, in, tile_sums, out, nullptr, tile, ntiles);
  grid_sync(bar);
}

template <class C>
__global__ void __launch_bounds__(SCAN_THREADS) msm_sort_small(SortArgs a) {
  const uint64_t gtid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x, gsize = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = gtid; i <= a.nbuckets; i += gsize) a.counts[i] = 0;
  for (uint64_t i = gtid; i < MSM_MAX_CHUNK + 2; i += gsize) a.hist[i] = 0;
  grid_sync(a.barrier);
  const uint64_t total = a.sh.n * a.sh.batch;
  for (uint64_t t = gtid; t < total; t += gsize) msm_digits_thread<C>(a.sh, a.scalars, a.digits, a.ranks, a.counts, t);
  grid_sync(a.barrier);
  sort_scan(a.barrier, a.nbuckets, 0, a.counts, a.tile_sums, a.starts);
  const uint64_t records = total * (uint64_t)a.sh.W;
  for (uint64_t t = gtid; t < records; t += gsize) msm_scatter_thread(a.sh, a.digits, a.starts, a.ranks, a.sorted, t);
  sort_scan(a.barrier, a.nbuckets, a.task_chunk, a.counts, a.tile_sums, a.task_first);
  msm_task_hist_body(a.nbuckets, a.task_chunk, a.starts, a.task_first, a.hist);
  grid_sync(a.barrier);
  msm_task_hist_scan_body(a.hist, a.hist_start, a.hist_cursor);
  grid_sync(a.barrier);
  msm_task_place_body(a.nbuckets, a.task_chunk, a.starts, a.task_first, a.hist_start, a.hist_cursor, a.tasks, a.slot_seg);
}

// ---------------------------------------------------------------------------------------------------------------
// Latency microbenchmarks behind kzg_gpu_microbench kinds 4..7 ("what bounds the tail"): the cost
// of one grid barrier, and of ONE dependent step of each kind with operands in registers and one warp per SM
// sub-partition — the numbers the step count of msm_tail has to be multiplied with.
static __global__ void __launch_bounds__(TAIL_THREADS, 1) bench_grid_barrier(int iters, GridBarrier* bar) {
  for (int i = 0; i < iters; i++) grid_sync(bar);
}
template <class F>
__global__ void __launch_bounds__(TAIL_THREADS, 1) bench_latency_chain(int kind, int iters, uint32_t seed, uint32_t* sink) {
  // a valid-looking (not necessarily on-curve) pair of projective points: the formulas do not care
  XYZZ<F> acc, p;
  F t = F::one();
  t.v[0] ^= (seed & 0xffu) + 3u;
  acc.X = F::mul(t, F::r2()); acc.Y = F::sqr(acc.X); acc.ZZ = F::mul(acc.Y, t); acc.ZZZ = F::sqr(acc.ZZ);
  p.X = F::sqr(acc.ZZZ); p.Y = F::mul(p.X, t); p.ZZ = F::sqr(p.Y); p.ZZZ = F::mul(p.ZZ, acc.X);
  F m = acc.X;
  for (int it = 0; it < iters; it++) {
    if (kind == 5) acc = team_add<F>(acc, p);   // 4 lanes per addition: 5 product rounds
    else if (kind == 6) acc.add(p);             // one thread per addition: 14 dependent products
    else if (kind == 8) m = F::inv(F::add(m, p.X));  // one field inversion (divsteps) per step
    else if (kind == 9) m = F::inv_bingcd(F::add(m, p.X));  // the binary GCD it replaced
    else m = F::mul(m, p.X);                    // kind 7: one dependent Montgomery product
  }
  uint32_t o = 0;
#pragma unroll
  for (int i = 0; i < Field<typename F::Params>::L; i++) o ^= acc.X.v[i] ^ acc.ZZZ.v[i] ^ m.v[i];
  if (o == 0x12345678u) sink[0] = o;
}

}  // namespace kzgb200
#endif
