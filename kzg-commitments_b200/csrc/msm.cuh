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
KZG_HD void msm_digits_thread(const MsmShape& sh, const uint8_t* scalars, int32_t* digits, uint32_t* counts, uint64_t t) {
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
      atomic_inc_u32(&counts[group * sh.nb + (mag - 1)]);
    }
  }
}
#if defined(__CUDACC__)
template <class C>
__global__ void msm_digits(MsmShape sh, const uint8_t* __restrict__ scalars, int32_t* __restrict__ digits,
                           uint32_t* __restrict__ counts) {
  const uint64_t total = sh.n * sh.batch;
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < total;
       t += (uint64_t)gridDim.x * blockDim.x)
    msm_digits_thread<C>(sh, scalars, digits, counts, t);
}
#endif

// ---------------------------------------------------------------- 2. scans (one CTA, decoupled by thread ranges)
// starts[b] = exclusive sum of counts; task_first[b] = exclusive sum of ceil(counts[b]/chunk).
// starts[nbuckets] / task_first[nbuckets] hold the totals.  cursor[] is cleared for the scatter.
#if defined(__CUDACC__)
static __global__ void msm_scan(uint64_t nbuckets, int chunk, const uint32_t* __restrict__ counts,
                         uint32_t* __restrict__ starts, uint32_t* __restrict__ task_first,
                         uint32_t* __restrict__ cursor) {
  __shared__ uint32_t s_cnt[1024], s_tsk[1024];
  const int T = blockDim.x, tid = threadIdx.x;
  const uint64_t per = (nbuckets + T - 1) / T;
  const uint64_t lo = min(per * tid, nbuckets), hi = min(lo + per, nbuckets);
  uint32_t a = 0, b = 0;
  for (uint64_t i = lo; i < hi; i++) {
    uint32_t cnt = counts[i];
    a += cnt;
    b += (cnt + chunk - 1) / chunk;
  }
  s_cnt[tid] = a;
  s_tsk[tid] = b;
  __syncthreads();
  if (tid == 0) {
    uint32_t ra = 0, rb = 0;
    for (int i = 0; i < T; i++) {
      uint32_t ta = s_cnt[i], tb = s_tsk[i];
      s_cnt[i] = ra; s_tsk[i] = rb;
      ra += ta; rb += tb;
    }
    starts[nbuckets] = ra;
    task_first[nbuckets] = rb;
  }
  __syncthreads();
  a = s_cnt[tid];
  b = s_tsk[tid];
  for (uint64_t i = lo; i < hi; i++) {
    uint32_t cnt = counts[i];
    starts[i] = a;
    task_first[i] = b;
    cursor[i] = 0;
    a += cnt;
    b += (cnt + chunk - 1) / chunk;
  }
}

// ---------------------------------------------------------------- 3. scatter (counting sort)
// record = source index | sign << 31
#endif
KZG_HD void msm_scatter_thread(const MsmShape& sh, const int32_t* digits, const uint32_t* starts, uint32_t* cursor,
                               uint32_t* sorted, uint64_t t) {
  const int32_t d = digits[t];
  if (d == 0) return;
  const uint64_t bw = t / sh.n, i = t - bw * sh.n;      // bw = blob*W + w
  const uint64_t blob = bw / sh.W;
  const uint32_t w = (uint32_t)(bw - blob * sh.W);
  const uint32_t mag = (uint32_t)(d < 0 ? -d : d);
  const uint64_t group = blob * sh.groups_per_msm + (sh.precomp ? 0 : w);
  const uint64_t key = group * sh.nb + (mag - 1);
  const uint32_t pos = starts[key] + atomic_inc_u32(&cursor[key]);
  const uint64_t src = sh.precomp ? ((uint64_t)w * sh.table_stride + sh.base_offset + i) : (sh.base_offset + i);
  sorted[pos] = (uint32_t)src | (d < 0 ? 0x80000000u : 0u);
}
#if defined(__CUDACC__)
static __global__ void msm_scatter(MsmShape sh, const int32_t* __restrict__ digits, const uint32_t* __restrict__ starts,
                            uint32_t* __restrict__ cursor, uint32_t* __restrict__ sorted) {
  const uint64_t total = sh.n * sh.batch * (uint64_t)sh.W;
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < total;
       t += (uint64_t)gridDim.x * blockDim.x)
    msm_scatter_thread(sh, digits, starts, cursor, sorted, t);
}
#endif

// ---------------------------------------------------------------- 4. bucket accumulation
// One thread per task.  A task is a contiguous slice (<= chunk records) of ONE bucket's records;
// the bucket owning task t is found by binary search in task_first[].
template <class F>
KZG_HD void msm_accumulate_thread(uint64_t nbuckets, const uint32_t* task_first, const uint32_t* starts,
                                  const uint32_t* sorted, const Affine<F>* bases, XYZZ<F>* partial, uint32_t t) {
  // largest b with task_first[b] <= t  (buckets with zero tasks share the value of their successor)
  uint64_t lo = 0, hi = nbuckets;  // invariant: task_first[lo] <= t < task_first[hi]
  while (hi - lo > 1) {
    uint64_t mid = (lo + hi) >> 1;
    if (task_first[mid] <= t) lo = mid; else hi = mid;
  }
  const uint64_t b = lo;
  const uint32_t k = t - task_first[b];
  const uint32_t nt = task_first[b + 1] - task_first[b];
  const uint32_t s0 = starts[b], cnt = starts[b + 1] - s0;
  const uint32_t e0 = s0 + (uint32_t)(((uint64_t)k * cnt) / nt);
  const uint32_t e1 = s0 + (uint32_t)(((uint64_t)(k + 1) * cnt) / nt);
  XYZZ<F> acc = XYZZ<F>::inf();
  for (uint32_t e = e0; e < e1; e++) {
    const uint32_t rec = sorted[e];
    const Affine<F> p = bases[rec & 0x7fffffffu];
    acc.madd(p, (rec >> 31) != 0);
  }
  partial[t] = acc;
}
#if defined(__CUDACC__)
template <class F>
__global__ void __launch_bounds__(128)
msm_accumulate(uint64_t nbuckets, const uint32_t* __restrict__ task_first, const uint32_t* __restrict__ starts,
               const uint32_t* __restrict__ sorted, const Affine<F>* __restrict__ bases,
               XYZZ<F>* __restrict__ partial) {
  const uint32_t ntasks = task_first[nbuckets];
  for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < ntasks; t += gridDim.x * blockDim.x)
    msm_accumulate_thread<F>(nbuckets, task_first, starts, sorted, bases, partial, t);
}
#endif

// ---------------------------------------------------------------- 5. bucket reduction
// chunk_out[group * chunks_per_group + j] = sum over buckets b in chunk j of (b+1) * B_b,
// B_b = sum of the bucket's task partials.  Running sums from the top of the chunk give
// sum (b-lo+1) B_b; the remaining lo * S is one short double-and-add.
template <class F>
KZG_HD void msm_reduce_chunk_thread(int nb, int buckets_per_chunk, const uint32_t* task_first,
                                    const XYZZ<F>* partial, XYZZ<F>* chunk_out, uint64_t t) {
  const int cpg = nb / buckets_per_chunk;
  const uint64_t g = t / cpg;
  const uint32_t j = (uint32_t)(t - g * cpg);
  const uint32_t lo = j * buckets_per_chunk;
  XYZZ<F> run = XYZZ<F>::inf(), acc = XYZZ<F>::inf();
  for (int b = buckets_per_chunk - 1; b >= 0; b--) {
    const uint64_t key = g * nb + lo + b;
    const uint32_t t0 = task_first[key], t1 = task_first[key + 1];
    for (uint32_t q = t0; q < t1; q++) run.add(partial[q]);
    acc.add(run);
  }
  if (lo != 0 && !run.is_inf()) {
    XYZZ<F> m = XYZZ<F>::inf();
    for (int bit = top_bit_u32(lo); bit >= 0; bit--) {
      m.dbl();
      if ((lo >> bit) & 1) m.add(run);
    }
    acc.add(m);
  }
  chunk_out[t] = acc;
}
#if defined(__CUDACC__)
template <class F>
__global__ void __launch_bounds__(128)
msm_reduce_chunks(uint64_t total_groups, int nb, int buckets_per_chunk, const uint32_t* __restrict__ task_first,
                  const XYZZ<F>* __restrict__ partial, XYZZ<F>* __restrict__ chunk_out) {
  const uint64_t total = total_groups * (uint64_t)(nb / buckets_per_chunk);
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < total;
       t += (uint64_t)gridDim.x * blockDim.x)
    msm_reduce_chunk_thread<F>(nb, buckets_per_chunk, task_first, partial, chunk_out, t);
}

// group_out[g] = sum of the group's chunk results.  One CTA per group, tree in shared memory.
template <class F, int THREADS>
__global__ void __launch_bounds__(THREADS)
msm_reduce_groups(int cpg, const XYZZ<F>* __restrict__ chunk_out, XYZZ<F>* __restrict__ group_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  XYZZ<F>* sm = reinterpret_cast<XYZZ<F>*>(smem_raw);
  const uint64_t g = blockIdx.x;
  XYZZ<F> acc = XYZZ<F>::inf();
  for (int j = threadIdx.x; j < cpg; j += THREADS) acc.add(chunk_out[g * cpg + j]);
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int s = THREADS / 2; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s) {
      acc.add(sm[threadIdx.x + s]);
      sm[threadIdx.x] = acc;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) group_out[g] = acc;
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
                              int modbytes, XYZZ<F>* out_xyzz, uint64_t m) {
  XYZZ<F> acc = group_out[m * groups_per_msm + (groups_per_msm - 1)];
  for (int g = groups_per_msm - 2; g >= 0; g--) {
    for (int k = 0; k < c; k++) acc.dbl();
    acc.add(group_out[m * groups_per_msm + g]);
  }
  if (out_xyzz) out_xyzz[m] = acc;
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
                           XYZZ<F>* __restrict__ out_xyzz) {
  const uint64_t m = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (m >= batch) return;
  msm_finish_thread<F>(groups_per_msm, c, group_out, out_bytes, is_inf, modbytes, out_xyzz, m);
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
