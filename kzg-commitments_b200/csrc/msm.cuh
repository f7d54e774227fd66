// Pippenger multi-scalar multiplication  sum_i c_i * G_i  for G1 (F = Fp) and G2 (F = Fp2).
//
// Replaces trusted_setup::polyeval_G1 / polyeval_G2 (src/trusted_setup.cpp:149-201), which run one
// GLV/GS scalar multiplication per term (PAIR_G1mul pair.cpp:876-924, PAIR_G2mul pair.cpp:927-983)
// and add the terms one by one.  The group element computed is the same; the schedule is:
//
//   1. msm_digits     scalars -> signed c-bit digits (d in [-2^(c-1), 2^(c-1)]), bucket histogram
//   2. scan           bucket sizes -> bucket offsets, and bucket -> task offsets (<= CH points/task)
//   3. msm_scatter    counting sort: (point index | sign) records grouped by bucket
//   4. msm_accumulate one thread per task: mixed-affine adds into an XYZZ accumulator
//   5. msm_reduce     per bucket group: sum_b (b+1) * B_b by chunked running sums + block tree
//   6. msm_finish     Horner over the window groups (c doublings per step), affine, big-endian bytes
//
// A "group" is a set of 2^(c-1) buckets.  Without precomputation there is one group per window;
// with a precomputed table  T[w][i] = 2^(c*w) * G_i  (built once per SRS, HBM-resident) all windows
// share ONE group per MSM and step 6 needs no doublings.  `batch` independent MSMs over the same
// bases (blob commitments) run in the same launches: the bucket key carries the blob index.
#pragma once
#include "ec.cuh"

namespace kzgb200 {

struct MsmShape {
  int c;                 // window bits
  int W;                 // windows per scalar
  int nb;                // buckets per group = 2^(c-1)
  int groups_per_msm;    // W, or 1 with a precomputed table
  int precomp;           // 1: sources come from T[w][i] at index w*table_stride + i
  uint64_t n;            // scalars per MSM
  uint64_t batch;        // independent MSMs
  uint64_t table_stride; // points per table row (= SRS size) when precomp
  uint64_t base_offset;  // first SRS point used
  int chunk;             // CH: max points per accumulate task
  uint64_t total_groups() const { return batch * (uint64_t)groups_per_msm; }
  uint64_t total_buckets() const { return total_groups() * (uint64_t)nb; }
};

// The per-thread bodies below are plain host/device functions so that the CPU tests can run the
// whole pipeline through the carry-flag emulation (csrc/host_check.cpp); the __global__ wrappers only
// add the grid-stride loop.
KZG_HD uint32_t atomic_inc_u32(uint32_t* p) {
#if defined(__CUDA_ARCH__)
  return atomicAdd(p, 1u);
#else
  return (*p)++;
#endif
}
KZG_HD void atomic_max_u32(uint32_t* p, uint32_t v) {
#if defined(__CUDA_ARCH__)
  atomicMax(p, v);
#else
  if (v > *p) *p = v;
#endif
}
KZG_HD int top_bit_u32(uint32_t v) {  // index of the highest set bit, v != 0
#if defined(__CUDA_ARCH__)
  return 31 - __clz(v);
#else
  return 31 - __builtin_clz(v);
#endif
}

// scalar (little-endian words, canonical or not) -> reduced below r by repeated subtraction
template <class C>
KZG_HD void load_scalar(const uint8_t* scalars, uint64_t idx, uint32_t* k) {
  const uint32_t* src = reinterpret_cast<const uint32_t*>(scalars + idx * C::SCALAR_BYTES);
  constexpr int SW = C::SCALAR_WORDS;
  constexpr int LR = C::FrP::L;
#pragma unroll
  for (int i = 0; i < LR; i++) k[i] = (i < SW) ? src[i] : 0u;
  // at most floor(2^(8*SB) / r) subtractions (6 for BN254, 2 for BLS12-381, 7 for BN158)
  for (int it = 0; it < 8; it++) {
    uint32_t t[LR];
    t[0] = ptx::sub_cc(k[0], C::FrP::mod(0));
#pragma unroll
    for (int i = 1; i < LR; i++) t[i] = ptx::subc_cc(k[i], C::FrP::mod(i));
    uint32_t borrow = ptx::subc(0, 0);
    if (borrow) break;
#pragma unroll
    for (int i = 0; i < LR; i++) k[i] = t[i];
  }
}

template <int LR>
KZG_HD uint32_t scalar_bits(const uint32_t* k, int pos, int width) {
  // bits [pos, pos+width) of a little-endian word array, width <= 24
  int w = pos >> 5, s = pos & 31;
  uint64_t lo = 0;
#pragma unroll
  for (int i = 0; i < LR; i++) {
    if (i == w) lo |= (uint64_t)k[i];
    if (i == w + 1) lo |= (uint64_t)k[i] << 32;
  }
  return (uint32_t)(lo >> s) & ((1u << width) - 1u);
}

// ---------------------------------------------------------------- 1. digits + histogram
// digits[(blob*W + w) * n + i] = signed digit ; counts[key]++ for non-zero digits
template <class C>
KZG_HD void msm_digits_thread(const MsmShape& sh, const uint8_t* scalars, int32_t* digits, uint32_t* ranks, uint32_t* counts,
                              uint64_t t) {
  const uint64_t blob = t / sh.n, i = t - blob * sh.n;
  uint32_t k[C::FrP::L];
  load_scalar<C>(scalars, t, k);
  uint32_t carry = 0;
  const uint32_t half = 1u << (sh.c - 1);
  for (int w = 0; w < sh.W; w++) {
    uint32_t raw = scalar_bits<C::FrP::L>(k, w * sh.c, sh.c) + carry;
    int32_t d;
    if (raw > half) { d = (int32_t)raw - (int32_t)(1u << sh.c); carry = 1; }
    else { d = (int32_t)raw; carry = 0; }
    digits[(blob * sh.W + w) * sh.n + i] = d;
    if (d != 0) {
      const uint32_t mag = (uint32_t)(d < 0 ? -d : d);
      const uint64_t group = blob * sh.groups_per_msm + (sh.precomp ? 0 : w);
      // the value returned by the histogram increment is this record's rank inside its bucket: the
      // scatter pass then needs no second round of atomics
      ranks[(blob * sh.W + w) * sh.n + i] = atomic_inc_u32(&counts[group * sh.nb + (mag - 1)]);
    }
  }
}
#if defined(__CUDACC__)
template <class C>
__global__ void msm_digits(MsmShape sh, const uint8_t* __restrict__ scalars, int32_t* __restrict__ digits,
                           uint32_t* __restrict__ ranks, uint32_t* __restrict__ counts, uint64_t t_begin,
                           uint64_t t_end) {
  // scalars [t_begin, t_end) of the n * batch: the host path launches one slice per arriving H2D chunk
  for (uint64_t t = t_begin + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < t_end;
       t += (uint64_t)gridDim.x * blockDim.x)
    msm_digits_thread<C>(sh, scalars, digits, ranks, counts, t);
}
#endif

// ---------------------------------------------------------------- 2. scans (multi-block)
// Exclusive prefix sums over the bucket table in three launches: per-tile sums, one CTA scanning the
// tile sums, per-tile rescan.  `chunk` != 0 scans ceil(x / chunk) instead of x (tasks per bucket).
// out[n] receives the total.
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;
KZG_HD uint32_t scan_value(uint32_t x, int chunk) { return chunk ? (x + (uint32_t)chunk - 1u) / (uint32_t)chunk : x; }
#if defined(__CUDACC__)
// The three phases as block-wide device bodies (tile = the block's tile index) so that the fused small-MSM sort
// kernel (msm_tail.cuh) can run them between grid barriers; the __global__ wrappers are the large-MSM path.
static __device__ __forceinline__ void scan_tile_sum_body(uint64_t n, int chunk, const uint32_t* __restrict__ in,
                                                           uint32_t* __restrict__ tile_sums, uint64_t tile) {
  __shared__ uint32_t sm[SCAN_THREADS / 32];
  const uint64_t base = tile * SCAN_TILE;
  uint32_t a = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    const uint64_t i = base + (uint64_t)k * SCAN_THREADS + threadIdx.x;
    if (i < n) a += scan_value(in[i], chunk);
  }
  for (int d = 16; d > 0; d >>= 1) a += __shfl_down_sync(0xffffffffu, a, d);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < SCAN_THREADS / 32; w++) t += sm[w];
    tile_sums[tile] = t;
  }
  __syncthreads();  // sm is reused by the caller's next tile
}
// one CTA (any power-of-two blockDim <= 1024): exclusive scan of tile_sums[0..ntiles) in place; total -> tile_sums[ntiles]
static __device__ __forceinline__ void scan_tiles_body(uint64_t ntiles, uint32_t* __restrict__ tile_sums) {
  __shared__ uint32_t s_part[1024];
  const int T = blockDim.x, tid = threadIdx.x;
  const uint64_t per = (ntiles + T - 1) / T;
  const uint64_t lo = min(per * (uint64_t)tid, ntiles), hi = min(lo + per, ntiles);
  uint32_t a = 0;
  for (uint64_t i = lo; i < hi; i++) a += tile_sums[i];
  s_part[tid] = a;
  __syncthreads();
  for (int d = 1; d < T; d <<= 1) {  // Hillis-Steele inclusive scan of the per-thread sums
    uint32_t v = tid >= d ? s_part[tid - d] : 0u;
    __syncthreads();
    s_part[tid] += v;
    __syncthreads();
  }
  uint32_t run = s_part[tid] - a;  // exclusive prefix of this thread's range
  for (uint64_t i = lo; i < hi; i++) {
    const uint32_t v = tile_sums[i];
    tile_sums[i] = run;
    run += v;
  }
  if (tid == T - 1) tile_sums[ntiles] = s_part[T - 1];
}
static __device__ __forceinline__ void scan_apply_body(uint64_t n, int chunk, const uint32_t* __restrict__ in,
                                                        const uint32_t* __restrict__ tile_sums, uint32_t* __restrict__ out,
                                                        uint32_t* __restrict__ zero_me, uint64_t tile, uint64_t ntiles) {
  __shared__ uint32_t sm[SCAN_THREADS];
  // thread t owns SCAN_ITEMS consecutive elements of the tile
  const uint64_t base = tile * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS], a = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    v[k] = (base + k < n) ? scan_value(in[base + k], chunk) : 0u;
    a += v[k];
  }
  sm[threadIdx.x] = a;
  __syncthreads();
  for (int d = 1; d < SCAN_THREADS; d <<= 1) {
    uint32_t u = (int)threadIdx.x >= d ? sm[threadIdx.x - d] : 0u;
    __syncthreads();
    sm[threadIdx.x] += u;
    __syncthreads();
  }
  uint32_t run = tile_sums[tile] + sm[threadIdx.x] - a;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    if (base + k < n) {
      out[base + k] = run;
      if (zero_me) zero_me[base + k] = 0;
    }
    run += v[k];
  }
  if (tile == ntiles - 1 && threadIdx.x == 0) out[n] = tile_sums[ntiles];
  __syncthreads();  // sm is reused by the caller's next tile
}
static __global__ void __launch_bounds__(SCAN_THREADS)
scan_tile_sums(uint64_t n, int chunk, const uint32_t* __restrict__ in, uint32_t* __restrict__ tile_sums) {
  scan_tile_sum_body(n, chunk, in, tile_sums, blockIdx.x);
}
static __global__ void __launch_bounds__(1024) scan_tiles(uint64_t ntiles, uint32_t* __restrict__ tile_sums) {
  scan_tiles_body(ntiles, tile_sums);
}
static __global__ void __launch_bounds__(SCAN_THREADS)
scan_apply(uint64_t n, int chunk, const uint32_t* __restrict__ in, const uint32_t* __restrict__ tile_sums,
           uint32_t* __restrict__ out, uint32_t* __restrict__ zero_me) {
  scan_apply_body(n, chunk, in, tile_sums, out, zero_me, blockIdx.x, gridDim.x);
}
#endif

// ---------------------------------------------------------------- 3. scatter (counting sort)
// record = source index | sign << 31
KZG_HD void msm_scatter_thread(const MsmShape& sh, const int32_t* digits, const uint32_t* starts, const uint32_t* ranks,
                               uint32_t* sorted, uint64_t t) {
  const int32_t d = digits[t];
  if (d == 0) return;
  const uint64_t bw = t / sh.n, i = t - bw * sh.n;      // bw = blob*W + w
  const uint64_t blob = bw / sh.W;
  const uint32_t w = (uint32_t)(bw - blob * sh.W);
  const uint32_t mag = (uint32_t)(d < 0 ? -d : d);
  const uint64_t group = blob * sh.groups_per_msm + (sh.precomp ? 0 : w);
  const uint64_t key = group * sh.nb + (mag - 1);
  const uint32_t pos = starts[key] + ranks[t];
  const uint64_t src = sh.precomp ? ((uint64_t)w * sh.table_stride + sh.base_offset + i) : (sh.base_offset + i);
  sorted[pos] = (uint32_t)src | (d < 0 ? 0x80000000u : 0u);
}
#if defined(__CUDACC__)
static __global__ void msm_scatter(MsmShape sh, const int32_t* __restrict__ digits, const uint32_t* __restrict__ starts,
                            const uint32_t* __restrict__ ranks, uint32_t* __restrict__ sorted) {
  const uint64_t total = sh.n * sh.batch * (uint64_t)sh.W;
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < total;
       t += (uint64_t)gridDim.x * blockDim.x)
    msm_scatter_thread(sh, digits, starts, ranks, sorted, t);
}
#endif

// ---------------------------------------------------------------- 4. tasks + bucket accumulation
// A task is a contiguous slice (<= chunk records) of ONE bucket's sorted records; bucket b owns tasks
// [task_first[b], task_first[b+1]) and its records are divided evenly among them.  Tasks are then
// ordered by DESCENDING length (counting sort on the length) so that the 32 lanes of a warp run the
// same trip count — bucket sizes are Poisson distributed and an unsorted assignment left 30 % of the
// lanes idle (profiles/r1_accumulate_v1.md) — and so that the longest tasks start first.
struct MsmTask {
  uint32_t first;  // index of the first record in sorted[]
  uint32_t len;    // number of records
  uint32_t slot;   // where the partial sum goes: partial[slot], slot = task_first[b] + k
};
struct MsmSlotSeg {
  uint32_t k, nt;  // partial[slot] is task k of a bucket with nt tasks (segment descriptor for the fold)
};
constexpr int MSM_MAX_CHUNK = 255;  // task-length histogram has MSM_MAX_CHUNK + 1 bins

KZG_HD void msm_task_of_bucket(const uint32_t* starts, const uint32_t* task_first, uint64_t b, uint32_t k, MsmTask& t) {
  const uint32_t nt = task_first[b + 1] - task_first[b];
  const uint32_t s0 = starts[b], cnt = starts[b + 1] - s0;
  const uint32_t e0 = s0 + (uint32_t)(((uint64_t)k * cnt) / nt);
  const uint32_t e1 = s0 + (uint32_t)(((uint64_t)(k + 1) * cnt) / nt);
  t.first = e0;
  t.len = e1 - e0;
  t.slot = task_first[b] + k;
}
// hist[chunk - len]++ for every task (bin 0 = longest)
KZG_HD void msm_task_hist_thread(int chunk, const uint32_t* starts, const uint32_t* task_first, uint32_t* hist, uint64_t b) {
  const uint32_t nt = task_first[b + 1] - task_first[b];
  if (nt > hist[MSM_MAX_CHUNK + 1]) atomic_max_u32(&hist[MSM_MAX_CHUNK + 1], nt);  // extra bin: most tasks in one bucket
  for (uint32_t k = 0; k < nt; k++) {
    MsmTask t;
    msm_task_of_bucket(starts, task_first, b, k, t);
    atomic_inc_u32(&hist[chunk - (int)t.len]);
  }
}
// tasks[hist_start[bin] + cursor[bin]++] = task
KZG_HD void msm_task_place_thread(int chunk, const uint32_t* starts, const uint32_t* task_first, const uint32_t* hist_start,
                                  uint32_t* cursor, MsmTask* tasks, MsmSlotSeg* slot_seg, uint64_t b) {
  const uint32_t nt = task_first[b + 1] - task_first[b];
  for (uint32_t k = 0; k < nt; k++) {
    MsmTask t;
    msm_task_of_bucket(starts, task_first, b, k, t);
    const int bin = chunk - (int)t.len;
    tasks[hist_start[bin] + atomic_inc_u32(&cursor[bin])] = t;
    slot_seg[t.slot] = MsmSlotSeg{k, nt};
  }
}
// point_stride: bytes from one source point to the next (sizeof(Affine<F>), or 128 for the padded window table of
// the "table_pad" experiment, profiles/r2_gather_layout.md)
template <class F>
KZG_HD void msm_accumulate_thread(const MsmTask* tasks, const uint32_t* sorted, const Affine<F>* bases,
                                  XYZZ<F>* partial, uint32_t i, size_t point_stride = sizeof(Affine<F>)) {
  const MsmTask t = tasks[i];
  XYZZ<F> acc = XYZZ<F>::inf();
  const char* base_bytes = reinterpret_cast<const char*>(bases);
  for (uint32_t e = t.first; e < t.first + t.len; e++) {
    const uint32_t rec = sorted ? sorted[e] : e;  // no record array: the points were pre-reduced (msm_affine.cuh)
    const Affine<F> p = *reinterpret_cast<const Affine<F>*>(base_bytes + (size_t)(rec & 0x7fffffffu) * point_stride);
    acc.madd(p, (rec >> 31) != 0);
  }
  partial[t.slot] = acc;
}
// Fold: partial[task_first[b]] = sum of bucket b's task partials, in place.  Afterwards the value of
// bucket b is partial[task_first[b]] if it has any task and infinity otherwise (bucket_value()).
// The device path (msm_fold_level) is a segmented binary tree; this per-thread body is the sequential
// statement of the same sum, used by the host emulation.
template <class F>
KZG_HD void msm_fold_thread(const uint32_t* task_first, XYZZ<F>* partial, uint64_t b) {
  const uint32_t t0 = task_first[b], t1 = task_first[b + 1];
  if (t1 - t0 < 2) return;
  XYZZ<F> acc = partial[t0];
  for (uint32_t q = t0 + 1; q < t1; q++) acc.add(partial[q]);
  partial[t0] = acc;
}
template <class F>
KZG_HD XYZZ<F> bucket_value(const uint32_t* task_first, const XYZZ<F>* partial, uint64_t b) {
  const uint32_t t0 = task_first[b];
  if (task_first[b + 1] == t0) return XYZZ<F>::inf();
  return partial[t0];
}
#if defined(__CUDACC__)
// task-length histogram, privatised in shared memory (block-wide body; grid-stride over the buckets)
static __device__ __forceinline__ void msm_task_hist_body(uint64_t nbuckets, int chunk, const uint32_t* __restrict__ starts,
                                                           const uint32_t* __restrict__ task_first, uint32_t* __restrict__ hist) {
  __shared__ uint32_t sh[MSM_MAX_CHUNK + 2];
  for (int i = threadIdx.x; i <= MSM_MAX_CHUNK + 1; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  for (uint64_t b = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < nbuckets; b += (uint64_t)gridDim.x * blockDim.x)
    msm_task_hist_thread(chunk, starts, task_first, sh, b);
  __syncthreads();
  for (int i = threadIdx.x; i <= MSM_MAX_CHUNK; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], sh[i]);
  if (threadIdx.x == 0 && sh[MSM_MAX_CHUNK + 1]) atomicMax(&hist[MSM_MAX_CHUNK + 1], sh[MSM_MAX_CHUNK + 1]);
}
static __global__ void __launch_bounds__(256)
msm_task_hist(uint64_t nbuckets, int chunk, const uint32_t* __restrict__ starts, const uint32_t* __restrict__ task_first,
              uint32_t* __restrict__ hist) {
  msm_task_hist_body(nbuckets, chunk, starts, task_first, hist);
}
// hist (MSM_MAX_CHUNK+1 bins) -> exclusive starts; cursors cleared.  One thread is plenty.
static __device__ __forceinline__ void msm_task_hist_scan_body(const uint32_t* __restrict__ hist, uint32_t* __restrict__ hist_start,
                                                                uint32_t* __restrict__ cursor) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  uint32_t run = 0;
  for (int i = 0; i <= MSM_MAX_CHUNK; i++) {
    hist_start[i] = run;
    run += hist[i];
    cursor[i] = 0;
  }
}
static __global__ void msm_task_hist_scan(const uint32_t* __restrict__ hist, uint32_t* __restrict__ hist_start,
                                          uint32_t* __restrict__ cursor) {
  msm_task_hist_scan_body(hist, hist_start, cursor);
}
// placement: a CTA first counts its tasks per bin in shared memory, reserves one range per bin with a
// single global atomic, then its threads take slots inside the reserved ranges.
static __device__ __forceinline__ void msm_task_place_body(uint64_t nbuckets, int chunk, const uint32_t* __restrict__ starts,
                                                            const uint32_t* __restrict__ task_first,
                                                            const uint32_t* __restrict__ hist_start, uint32_t* __restrict__ cursor,
                                                            MsmTask* __restrict__ tasks, MsmSlotSeg* __restrict__ slot_seg) {
  __shared__ uint32_t cnt[MSM_MAX_CHUNK + 2], base[MSM_MAX_CHUNK + 1];
  for (uint64_t b0 = (uint64_t)blockIdx.x * blockDim.x; b0 < nbuckets; b0 += (uint64_t)gridDim.x * blockDim.x) {
    for (int i = threadIdx.x; i <= MSM_MAX_CHUNK + 1; i += blockDim.x) cnt[i] = 0;
    __syncthreads();
    const uint64_t b = b0 + threadIdx.x;
    if (b < nbuckets) msm_task_hist_thread(chunk, starts, task_first, cnt, b);
    __syncthreads();
    for (int i = threadIdx.x; i <= MSM_MAX_CHUNK; i += blockDim.x) {
      base[i] = cnt[i] ? hist_start[i] + atomicAdd(&cursor[i], cnt[i]) : 0u;
      cnt[i] = 0;
    }
    __syncthreads();
    if (b < nbuckets) msm_task_place_thread(chunk, starts, task_first, base, cnt, tasks, slot_seg, b);
    __syncthreads();
  }
}
static __global__ void __launch_bounds__(256)
msm_task_place(uint64_t nbuckets, int chunk, const uint32_t* __restrict__ starts, const uint32_t* __restrict__ task_first,
               const uint32_t* __restrict__ hist_start, uint32_t* __restrict__ cursor, MsmTask* __restrict__ tasks,
               MsmSlotSeg* __restrict__ slot_seg) {
  msm_task_place_body(nbuckets, chunk, starts, task_first, hist_start, cursor, tasks, slot_seg);
}
template <class F>
__global__ void __launch_bounds__(128, sizeof(F) <= 48 ? 4 : 1)
msm_accumulate(const uint32_t* __restrict__ ntasks_ptr, const MsmTask* __restrict__ tasks,
               const uint32_t* __restrict__ sorted, const Affine<F>* __restrict__ bases, XYZZ<F>* __restrict__ partial,
               size_t point_stride) {
  const uint32_t ntasks = *ntasks_ptr;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ntasks; i += gridDim.x * blockDim.x)
    msm_accumulate_thread<F>(tasks, sorted, bases, partial, i, point_stride);
}
// Device fold: a segmented binary tree over the flat array of task partials, one launch per level.
// At level l the partial of task k (k a multiple of 2^(l+1)) absorbs that of task k + 2^l of the same
// bucket, so a bucket with nt tasks is folded in ceil(log2 nt) single additions instead of nt - 1
// sequential ones (a sequential XYZZ addition chain runs at ~19 us per link; with structured scalars
// or a short top window a few buckets own hundreds of tasks).  Levels beyond the largest task count
// (max_nt, found by msm_task_hist) return immediately.
template <class F>
__global__ void __launch_bounds__(128)
msm_fold_level(int level, const uint32_t* __restrict__ ntasks_ptr, const uint32_t* __restrict__ max_nt,
               const MsmSlotSeg* __restrict__ slot_seg, XYZZ<F>* __restrict__ partial) {
  const uint32_t stride = 1u << level;
  if (stride >= *max_nt) return;
  const uint32_t ntasks = *ntasks_ptr;
  for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < ntasks; q += gridDim.x * blockDim.x) {
    const MsmSlotSeg sg = slot_seg[q];
    if ((sg.k & (2u * stride - 1u)) == 0u && sg.k + stride < sg.nt) partial[q].add(partial[q + stride]);
  }
}
#endif

// ---------------------------------------------------------------- 5. bucket reduction
// Per group of nb = R * Cn buckets (Cn = 2^h):  sum_b (b+1) B_b  with b = hi*Cn + lo  equals
//     Cn * sum_hi hi * row[hi]  +  sum_lo (lo+1) * col[lo],
// row[hi] = sum_lo B[hi][lo], col[lo] = sum_hi B[hi][lo].  Every step below is a single group
// operation per thread, because one thread's XYZZ addition is a ~19 us dependent carry chain
// (profiles/r1b): the 2*nb unweighted additions of the row/column sums are BINARY trees (one launch
// per level, rows and columns side by side in blockIdx.y), and the two short weighted sums run in
// chunks of WCHUNK terms (running sums + a double-and-add correction for the chunk offset) followed
// by the same pairwise trees.  All lengths are powers of two.
constexpr int WCHUNK = 2;  // terms per thread of the weighted sums

// level 0 of a row/column tree, reading bucket values through task_first:
// out[(g*nsums + s)*P + p] = B[g*nb + s*sstride + 2p*jstride] + B[.. + (2p+1)*jstride],  P = max(len/2, 1)
template <class F>
struct BucketJob {
  XYZZ<F>* out;
  uint64_t nsums, len, sstride, jstride, total;  // total = groups * nsums * P
};
template <class F>
KZG_HD void msm_bucket_pair_thread(const BucketJob<F>& jb, uint64_t nb, const uint32_t* task_first, const XYZZ<F>* partial,
                                   uint64_t t) {
  const uint64_t P = jb.len >= 2 ? jb.len / 2 : 1;
  const uint64_t p = t % P, gs = t / P;
  const uint64_t s = gs % jb.nsums, g = gs / jb.nsums;
  const uint64_t b0 = g * nb + s * jb.sstride + 2 * p * jb.jstride;
  XYZZ<F> acc = bucket_value<F>(task_first, partial, b0);
  if (jb.len >= 2) acc.add(bucket_value<F>(task_first, partial, b0 + jb.jstride));
  jb.out[t] = acc;
}
// further levels: out[i] = in[2i] + in[2i+1]
template <class F>
struct PairJob {
  const XYZZ<F>* in;
  XYZZ<F>* out;
  uint64_t n_out;
};
template <class F>
KZG_HD void msm_pair_sum_thread(const PairJob<F>& jb, uint64_t i) {
  XYZZ<F> acc = jb.in[2 * i];
  acc.add(jb.in[2 * i + 1]);
  jb.out[i] = acc;
}
// k * P for a small non-negative k (chunk offsets of the weighted sums)
template <class F>
KZG_HD XYZZ<F> small_mul(const XYZZ<F>& p, uint32_t k) {
  XYZZ<F> m = XYZZ<F>::inf();
  if (k == 0 || p.is_inf()) return m;
  for (int bit = top_bit_u32(k); bit >= 0; bit--) {
    m.dbl();
    if ((k >> bit) & 1) m.add(p);
  }
  return m;
}
// out[g*P + p] = sum_{j in chunk p} (j + woff) * in[g*m + j],  chunk p = [p*WCHUNK, ...), P = max(m/WCHUNK, 1)
template <class F>
struct WeightedJob {
  const XYZZ<F>* in;
  XYZZ<F>* out;
  uint64_t m, total;  // total = groups * P
  uint32_t woff;
};
template <class F>
KZG_HD void msm_weighted_chunk_thread(const WeightedJob<F>& jb, uint64_t t) {
  const uint64_t m = jb.m;
  const uint64_t P = m >= WCHUNK ? m / WCHUNK : 1;
  const uint64_t p = t % P, g = t / P;
  const uint64_t j0 = p * WCHUNK, j1 = (j0 + WCHUNK < m) ? j0 + WCHUNK : m;
  XYZZ<F> run = XYZZ<F>::inf(), acc = XYZZ<F>::inf();
  for (uint64_t j = j1; j > j0; j--) {  // acc = sum (j - j0 + 1) * in[j]
    run.add(jb.in[g * m + j - 1]);
    acc.add(run);
  }
  const uint32_t k = (uint32_t)j0 + jb.woff;  // weights wanted: j + woff = (j - j0 + 1) + (j0 + woff - 1)
  if (k >= 1) acc.add(small_mul<F>(run, k - 1));
  else {       // k == 0 (rows, first chunk): weights j - j0, i.e. one `run` too many
    run.negate();
    acc.add(run);
  }
  jb.out[t] = acc;
}
// group_out[g] = 2^h * T1[g] + T2[g]
template <class F>
KZG_HD void msm_group_combine_thread(int h, const XYZZ<F>* t1, const XYZZ<F>* t2, XYZZ<F>* group_out, uint64_t g) {
  XYZZ<F> acc = t1[g];
  for (int k = 0; k < h; k++) acc.dbl();
  acc.add(t2[g]);
  group_out[g] = acc;
}
#if defined(__CUDACC__)
template <class F>
__global__ void __launch_bounds__(128)
msm_bucket_pairs(uint64_t nb, const uint32_t* __restrict__ task_first, const XYZZ<F>* __restrict__ partial,
                 BucketJob<F> j0, BucketJob<F> j1) {
  const BucketJob<F>& jb = blockIdx.y ? j1 : j0;
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < jb.total; t += (uint64_t)gridDim.x * blockDim.x)
    msm_bucket_pair_thread<F>(jb, nb, task_first, partial, t);
}
template <class F>
__global__ void __launch_bounds__(128) msm_pair_sum(PairJob<F> j0, PairJob<F> j1) {
  const PairJob<F>& jb = blockIdx.y ? j1 : j0;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < jb.n_out; i += (uint64_t)gridDim.x * blockDim.x)
    msm_pair_sum_thread<F>(jb, i);
}
template <class F>
__global__ void __launch_bounds__(128) msm_weighted_chunks(WeightedJob<F> j0, WeightedJob<F> j1) {
  const WeightedJob<F>& jb = blockIdx.y ? j1 : j0;
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < jb.total; t += (uint64_t)gridDim.x * blockDim.x)
    msm_weighted_chunk_thread<F>(jb, t);
}
template <class F>
__global__ void __launch_bounds__(128)
msm_group_combine(uint64_t groups, int h, const XYZZ<F>* __restrict__ t1, const XYZZ<F>* __restrict__ t2,
                  XYZZ<F>* __restrict__ group_out) {
  const uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (g < groups) msm_group_combine_thread<F>(h, t1, t2, group_out, g);
}
#endif

// big-endian canonical bytes of a field element (MODBYTES each), as FP_toBytes / BIG_toBytes write them
template <class P>
KZG_HDN void fe_to_be_bytes(const Field<P>& mont, uint8_t* out, int modbytes) {
  Field<P> c = Field<P>::from_mont(mont);
  for (int j = 0; j < modbytes; j++) {
    int bytepos = modbytes - 1 - j;  // little-endian byte index
    out[j] = (uint8_t)(c.v[bytepos >> 2] >> (8 * (bytepos & 3)));
  }
}
template <class P>
KZG_HDN Field<P> fe_from_be_bytes(const uint8_t* in, int modbytes) {
  Field<P> c = Field<P>::zero();
  for (int j = 0; j < modbytes; j++) {
    int bytepos = modbytes - 1 - j;
    c.v[bytepos >> 2] |= (uint32_t)in[j] << (8 * (bytepos & 3));
  }
  return Field<P>::to_mont(c);
}

// G1: x | y ; G2: x.b | x.a | y.b | y.a  (imaginary part first, fp2.cpp:83-87, ecp2.cpp:209-215)
template <class P>
KZG_HD void affine_to_bytes(const Affine<Field<P>>& p, uint8_t* out, int M) {
  fe_to_be_bytes<P>(p.x, out, M);
  fe_to_be_bytes<P>(p.y, out + M, M);
}
template <class P>
KZG_HD void affine_to_bytes(const Affine<Field2<P>>& p, uint8_t* out, int M) {
  fe_to_be_bytes<P>(p.x.b, out, M);
  fe_to_be_bytes<P>(p.x.a, out + M, M);
  fe_to_be_bytes<P>(p.y.b, out + 2 * M, M);
  fe_to_be_bytes<P>(p.y.a, out + 3 * M, M);
}
template <class P>
KZG_HD void affine_from_bytes(Affine<Field<P>>& p, const uint8_t* in, int M) {
  p.x = fe_from_be_bytes<P>(in, M);
  p.y = fe_from_be_bytes<P>(in + M, M);
}
template <class P>
KZG_HD void affine_from_bytes(Affine<Field2<P>>& p, const uint8_t* in, int M) {
  p.x.b = fe_from_be_bytes<P>(in, M);
  p.x.a = fe_from_be_bytes<P>(in + M, M);
  p.y.b = fe_from_be_bytes<P>(in + 2 * M, M);
  p.y.a = fe_from_be_bytes<P>(in + 3 * M, M);
}
template <class F> struct CoordCount;
template <class P> struct CoordCount<Field<P>> { static constexpr int value = 2; };
template <class P> struct CoordCount<Field2<P>> { static constexpr int value = 4; };

// ---------------------------------------------------------------- 6. finish
// One thread per MSM of the batch: Horner over its window groups, optional extra terms, affine, bytes.
// out_xyzz (optional) also receives the projective result (used by the multi-GPU partial sums).
template <class F>
KZG_HD void msm_finish_thread(int groups_per_msm, int c, const XYZZ<F>* group_out, uint8_t* out_bytes, int* is_inf,
                              int modbytes, XYZZ<F>* out_xyzz, uint64_t m, int want_affine = 1) {
  XYZZ<F> acc = group_out[m * groups_per_msm + (groups_per_msm - 1)];
  for (int g = groups_per_msm - 2; g >= 0; g--) {
    for (int k = 0; k < c; k++) acc.dbl();
    acc.add(group_out[m * groups_per_msm + g]);
  }
  if (out_xyzz) out_xyzz[m] = acc;
  if (!want_affine) {  // the caller only wants the projective partial (point-range split across GPUs)
    is_inf[m] = acc.is_inf() ? 1 : 0;
    return;
  }
  const int nbytes = CoordCount<F>::value * modbytes;
  if (acc.is_inf()) {
    for (int j = 0; j < nbytes; j++) out_bytes[m * nbytes + j] = 0;
    is_inf[m] = 1;
  } else {
    affine_to_bytes(acc.to_affine(), out_bytes + m * nbytes, modbytes);
    is_inf[m] = 0;
  }
}
#if defined(__CUDACC__)
template <class F>
__global__ void msm_finish(uint64_t batch, int groups_per_msm, int c, const XYZZ<F>* __restrict__ group_out,
                           uint8_t* __restrict__ out_bytes, int* __restrict__ is_inf, int modbytes,
                           XYZZ<F>* __restrict__ out_xyzz, int want_affine, int lane_stride) {
  // lane_stride = 32: one MSM per WARP (lane 0 works).  The binary-GCD inversion has data-dependent loops; 32
  // different inversions in one warp diverge and run ~3x longer than one (0.33 ms against 0.10 ms measured).
  const uint64_t gt = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (gt % (uint64_t)lane_stride) return;
  const uint64_t m = gt / (uint64_t)lane_stride;
  if (m >= batch) return;
  msm_finish_thread<F>(groups_per_msm, c, group_out, out_bytes, is_inf, modbytes, out_xyzz, m, want_affine);
}

// out = sum of n projective points (the fold of the per-GPU partial sums of a point-range split)
template <class F>
__global__ void sum_points(uint64_t n, const XYZZ<F>* __restrict__ pts, uint8_t* __restrict__ out_bytes,
                           int* __restrict__ is_inf, int modbytes) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  XYZZ<F> acc = XYZZ<F>::inf();
  for (uint64_t i = 0; i < n; i++) acc.add(pts[i]);
  const int nbytes = CoordCount<F>::value * modbytes;
  if (acc.is_inf()) {
    for (int j = 0; j < nbytes; j++) out_bytes[j] = 0;
    is_inf[0] = 1;
  } else {
    affine_to_bytes(acc.to_affine(), out_bytes, modbytes);
    is_inf[0] = 0;
  }
}

// SRS bytes -> Montgomery affine points (one thread per point)
template <class F>
__global__ void points_from_bytes(uint64_t n, const uint8_t* __restrict__ in, int modbytes, Affine<F>* __restrict__ out) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  Affine<F> p;
  affine_from_bytes(p, in + i * (uint64_t)(CoordCount<F>::value * modbytes), modbytes);
  out[i] = p;
}
#endif  // __CUDACC__

}  // namespace kzgb200
