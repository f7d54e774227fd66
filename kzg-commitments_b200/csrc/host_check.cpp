// Host build of the field / group-law / MSM-pipeline headers (g++, no CUDA): the PTX carry-chain
// primitives run through the emulation in ptx.cuh, the per-thread kernel bodies of msm.cuh are driven
// by plain loops.  Loaded by the CPU tests (tests/test_host_emulation.py) through ctypes so that limb
// arithmetic, exceptional group-law cases, signed-digit extraction, bucket weights and the reduction
// tree are checked against the oracle on the CPU box before any GPU time is spent.
// NOT part of the product path: nothing in libkzg_gpu.so or the Python mirror calls into this file.
#include <cstdint>
#include <cstring>
#include <vector>

#include "curve_constants.cuh"
#include "diag.cuh"
#include "msm.cuh"
#include "msm_affine.cuh"
#include "msm_direct.cuh"

using namespace kzgb200;

namespace {

template <class F>
int field_op(int op, const uint32_t* a, const uint32_t* b, uint32_t* out) {
  F x, y;
  for (int i = 0; i < F::L; i++) { x.v[i] = a[i]; y.v[i] = b[i]; }
  x = F::to_mont(x);
  y = F::to_mont(y);
  F r;
  switch (op) {
    case 0: r = F::mul(x, y); break;
    case 1: r = F::add(x, y); break;
    case 2: r = F::sub(x, y); break;
    case 3: r = F::inv(x); break;
    case 4: r = F::neg(x); break;
    default: return -1;
  }
  r = F::from_mont(r);
  for (int i = 0; i < F::L; i++) out[i] = r.v[i];
  return 0;
}

// The Pippenger pipeline of ctx.cuh::enqueue_msm with loops in place of launches.
template <class C, class F>
int msm_host(uint64_t srs_n, const uint8_t* pts_bytes, const uint8_t* scalars, uint64_t count, uint64_t batch,
             uint64_t base_offset, int c, int chunk, int bpc, int precomp, uint8_t* out, int* is_inf) {
  const int M = C::MODBYTES;
  std::vector<Affine<F>> bases(srs_n);
  for (uint64_t i = 0; i < srs_n; i++) affine_from_bytes(bases[i], pts_bytes + i * (uint64_t)(CoordCount<F>::value * M), M);
  MsmShape sh;
  sh.c = c;
  sh.W = (C::SCALAR_BITS + 1 + c - 1) / c;
  sh.nb = 1 << (c - 1);
  sh.precomp = precomp;
  sh.groups_per_msm = precomp ? 1 : sh.W;
  sh.n = count;
  sh.batch = batch;
  sh.table_stride = srs_n;
  sh.base_offset = base_offset;
  sh.chunk = chunk;
  std::vector<Affine<F>> table;
  if (precomp) {  // T[w][i] = 2^(c*w) * G_i  (what msm_precompute_table builds on device)
    table.resize((size_t)sh.W * srs_n);
    for (uint64_t i = 0; i < srs_n; i++) {
      XYZZ<F> cur = XYZZ<F>::from_affine(bases[i]);
      table[i] = bases[i];
      for (int w = 1; w < sh.W; w++) {
        for (int k = 0; k < c; k++) cur.dbl();
        table[(size_t)w * srs_n + i] = cur.to_affine();
      }
    }
  }
  const uint64_t records = count * batch * sh.W;
  const uint64_t nbuckets = sh.total_buckets();
  std::vector<int32_t> digits(records);
  std::vector<uint32_t> ranks(records);
  std::vector<uint32_t> counts(nbuckets + 1, 0), starts(nbuckets + 1), task_first(nbuckets + 1), cursor(nbuckets + 1, 0),
      sorted(records);
  for (uint64_t t = 0; t < count * batch; t++) msm_digits_thread<C>(sh, scalars, digits.data(), ranks.data(), counts.data(), t);
  const int levels = (CoordCount<F>::value == 2) ? (bpc & 3) : 0;  // affine pre-reduction levels (G1 only)
  uint32_t a = 0, b = 0;
  for (uint64_t i = 0; i < nbuckets; i++) {
    starts[i] = a; task_first[i] = b;
    a += counts[i]; b += scan_value(counts[i], chunk << levels);
  }
  starts[nbuckets] = a; task_first[nbuckets] = b;
  for (uint64_t t = 0; t < records; t++) msm_scatter_thread(sh, digits.data(), starts.data(), ranks.data(), sorted.data(), t);
  // batched-affine pre-reduction (msm_affine.cuh), loops in place of launches
  const uint32_t* seg_off = starts.data();
  const uint32_t* seg_sorted = sorted.data();
  const Affine<F>* seg_pts = precomp ? table.data() : bases.data();
  std::vector<std::vector<uint32_t>> offs(levels);
  std::vector<std::vector<Affine<F>>> lvl_pts(levels);
  for (int l = 1; l <= levels; l++) {
    std::vector<uint32_t>& off = offs[l - 1];
    off.resize(nbuckets + 1);
    uint32_t run = 0;
    for (uint64_t i = 0; i < nbuckets; i++) { off[i] = run; run += scan_value(counts[i], 1 << l); }
    off[nbuckets] = run;
    lvl_pts[l - 1].resize(run + 1);
    std::vector<F> prefix(run + 1), totals(run / AFF_K + 2);
    AffLevel<F> lv{seg_sorted, seg_pts, seg_off, off.data(), nbuckets, lvl_pts[l - 1].data(), prefix.data(), totals.data()};
    const uint64_t nthreads = ((uint64_t)run + AFF_K - 1) / AFF_K;
    for (uint64_t t = 0; t < nthreads; t++) aff_forward_thread<F>(lv, t);
    for (uint64_t t = 0; t < nthreads; t++) totals[t] = F::inv(totals[t]);  // device: aff_invert (shared inversions)
    for (uint64_t t = 0; t < nthreads; t++) aff_backward_thread<F>(lv, t);
    seg_sorted = nullptr;
    seg_pts = lvl_pts[l - 1].data();
    seg_off = off.data();
  }
  // tasks ordered by descending length (msm_task_hist / msm_task_place)
  std::vector<uint32_t> hist(MSM_MAX_CHUNK + 2, 0), hist_start(MSM_MAX_CHUNK + 1), hcur(MSM_MAX_CHUNK + 1, 0);
  for (uint64_t i = 0; i < nbuckets; i++) msm_task_hist_thread(chunk, seg_off, task_first.data(), hist.data(), i);
  uint32_t run = 0;
  for (int i = 0; i <= MSM_MAX_CHUNK; i++) { hist_start[i] = run; run += hist[i]; }
  if (run != b) return -2;
  std::vector<MsmTask> tasks(b + 1);
  std::vector<MsmSlotSeg> slot_seg(b + 1);
  for (uint64_t i = 0; i < nbuckets; i++)
    msm_task_place_thread(chunk, seg_off, task_first.data(), hist_start.data(), hcur.data(), tasks.data(), slot_seg.data(), i);
  for (uint32_t t = 1; t < b; t++) if (tasks[t].len > tasks[t - 1].len) return -3;
  std::vector<XYZZ<F>> partial(b + 1);
  for (uint32_t t = 0; t < b; t++)
    msm_accumulate_thread<F>(tasks.data(), seg_sorted, seg_pts, partial.data(), t);
  for (uint64_t i = 0; i < nbuckets; i++) msm_fold_thread<F>(task_first.data(), partial.data(), i);
  // 2-D bucket reduction (ctx.cuh::enqueue_msm with loops in place of launches)
  const uint64_t groups = sh.total_groups();
  const int h = (sh.c - 1) / 2;
  const uint64_t Cn = 1ull << h, R = (uint64_t)sh.nb / Cn;
  auto tree = [&](std::vector<XYZZ<F>> in, uint64_t sums, uint64_t len) {
    while (len > 1) {
      std::vector<XYZZ<F>> out(sums * (len / 2));
      PairJob<F> jb{in.data(), out.data(), sums * (len / 2)};
      for (uint64_t i = 0; i < jb.n_out; i++) msm_pair_sum_thread<F>(jb, i);
      in.swap(out);
      len /= 2;
    }
    return in;
  };
  const uint64_t Lr = Cn >= 2 ? Cn / 2 : 1, Lc = R >= 2 ? R / 2 : 1;
  std::vector<XYZZ<F>> r0(groups * R * Lr), c0(groups * Cn * Lc);
  BucketJob<F> jr{r0.data(), R, Cn, Cn, 1, groups * R * Lr}, jc{c0.data(), Cn, R, 1, Cn, groups * Cn * Lc};
  for (uint64_t t = 0; t < jr.total; t++) msm_bucket_pair_thread<F>(jr, sh.nb, task_first.data(), partial.data(), t);
  for (uint64_t t = 0; t < jc.total; t++) msm_bucket_pair_thread<F>(jc, sh.nb, task_first.data(), partial.data(), t);
  std::vector<XYZZ<F>> rows = tree(r0, groups * R, Lr), cols = tree(c0, groups * Cn, Lc);
  const uint64_t Pr = R >= WCHUNK ? R / WCHUNK : 1, Pc = Cn >= WCHUNK ? Cn / WCHUNK : 1;
  std::vector<XYZZ<F>> wr(groups * Pr), wc(groups * Pc), group_out(groups);
  WeightedJob<F> jwr{rows.data(), wr.data(), R, groups * Pr, 0u}, jwc{cols.data(), wc.data(), Cn, groups * Pc, 1u};
  for (uint64_t t = 0; t < jwr.total; t++) msm_weighted_chunk_thread<F>(jwr, t);
  for (uint64_t t = 0; t < jwc.total; t++) msm_weighted_chunk_thread<F>(jwc, t);
  std::vector<XYZZ<F>> t1 = tree(wr, groups, Pr), t2 = tree(wc, groups, Pc);
  for (uint64_t g = 0; g < groups; g++) msm_group_combine_thread<F>(h, t1.data(), t2.data(), group_out.data(), g);
  for (uint64_t m = 0; m < batch; m++)
    msm_finish_thread<F>(sh.groups_per_msm, sh.c, group_out.data(), out, is_inf, M, (XYZZ<F>*)nullptr, m);
  return 0;
}

// The direct-lookup path of ctx.cuh::enqueue_msm_direct (msm_direct.cuh) with loops in place of launches: the table of
// every digit multiple, one partial sum per (scalars, window group) slot, pairwise trees, normalisation.  G1 only.
template <class C>
int msm_direct_host(uint64_t srs_n, const uint8_t* pts_bytes, const uint8_t* scalars, uint64_t count, uint64_t batch,
                    uint64_t base_offset, int c, int wpt, int spt, uint8_t* out, int* is_inf) {
  using F = Field<typename C::FpP>;
  const int M = C::MODBYTES;
  std::vector<Affine<F>> bases(srs_n);
  for (uint64_t i = 0; i < srs_n; i++) affine_from_bytes(bases[i], pts_bytes + i * (uint64_t)(2 * M), M);
  DirectShape sh{};
  sh.c = c;
  sh.W = (C::SCALAR_BITS + 1 + c - 1) / c;
  sh.H = 1ull << (c - 1);
  sh.n_table = srs_n;
  sh.n = count;
  sh.batch = batch;
  sh.base_offset = base_offset;
  sh.wpt = wpt;
  sh.groups = (sh.W + wpt - 1) / wpt;
  sh.spt = sh.groups > 1 ? 1 : spt;
  uint64_t Q = (count + sh.spt - 1) / sh.spt;
  if (sh.groups > 1) Q = (Q + 31) / 32 * 32;
  sh.P = Q * (uint64_t)sh.groups;
  sh.P2 = 1;
  while (sh.P2 < sh.P) sh.P2 <<= 1;
  // D[w][i][0] = 2^(c w) G_i (device: msm_precompute_table with a row of H entries as its stride), then the fill levels
  const uint64_t rows = (uint64_t)sh.W * srs_n;
  std::vector<Affine<F>> direct(rows * sh.H);
  for (uint64_t i = 0; i < srs_n; i++) {
    XYZZ<F> cur = XYZZ<F>::from_affine(bases[i]);
    direct[i * sh.H] = bases[i];
    for (int w = 1; w < sh.W; w++) {
      for (int k = 0; k < c; k++) cur.dbl();
      direct[((uint64_t)w * srs_n + i) * sh.H] = cur.is_inf() ? Affine<F>::inf() : cur.to_affine();
    }
  }
  for (int level = 0; level + 1 < c; level++)
    for (uint64_t t = 0; t < (rows << level); t++) direct_fill_thread<F>(level, sh.H, direct.data(), t);
  std::vector<XYZZ<F>> in(batch * sh.P2);
  for (uint64_t t = 0; t < batch * sh.P2; t++) direct_accumulate_thread<C, F>(sh, scalars, direct.data(), in.data(), t);
  for (uint64_t len = sh.P2; len > 1; len /= 2) {
    std::vector<XYZZ<F>> nxt(batch * (len / 2));
    PairJob<F> jb{in.data(), nxt.data(), batch * (len / 2)};
    for (uint64_t i = 0; i < jb.n_out; i++) msm_pair_sum_thread<F>(jb, i);
    in.swap(nxt);
  }
  for (uint64_t m = 0; m < batch; m++) msm_finish_thread<F>(1, c, in.data(), out, is_inf, M, (XYZZ<F>*)nullptr, m);
  return 0;
}

template <class C>
int dispatch_field(int which, int op, const uint32_t* a, const uint32_t* b, uint32_t* out) {
  return which == 0 ? field_op<Field<typename C::FpP>>(op, a, b, out) : field_op<Field<typename C::FrP>>(op, a, b, out);
}
template <class C>
int dispatch_msm(int group, uint64_t srs_n, const uint8_t* pts, const uint8_t* scalars, uint64_t count, uint64_t batch,
                 uint64_t base_offset, int c, int chunk, int bpc, int precomp, uint8_t* out, int* is_inf) {
  if (group == 1)
    return msm_host<C, Field<typename C::FpP>>(srs_n, pts, scalars, count, batch, base_offset, c, chunk, bpc, precomp, out, is_inf);
  return msm_host<C, Field2<typename C::FpP>>(srs_n, pts, scalars, count, batch, base_offset, c, chunk, bpc, precomp, out, is_inf);
}

}  // namespace

extern "C" {

int hc_selftest(int curve, uint32_t seed) {
  switch (curve) {
    case 0: return selftest_one<Curve_BN158>(seed);
    case 1: return selftest_one<Curve_BN254>(seed);
    case 2: return selftest_one<Curve_BLS12381>(seed);
  }
  return -1;
}

int hc_limbs(int curve, int which) {
  switch (curve) {
    case 0: return which == 0 ? Fp_BN158::L : Fr_BN158::L;
    case 1: return which == 0 ? Fp_BN254::L : Fr_BN254::L;
    case 2: return which == 0 ? Fp_BLS12381::L : Fr_BLS12381::L;
  }
  return -1;
}

// canonical little-endian limb arrays in and out; which: 0 = Fp, 1 = Fr; op: 0 mul 1 add 2 sub 3 inv 4 neg
int hc_field_op(int curve, int which, int op, const uint32_t* a, const uint32_t* b, uint32_t* out) {
  switch (curve) {
    case 0: return dispatch_field<Curve_BN158>(which, op, a, b, out);
    case 1: return dispatch_field<Curve_BN254>(which, op, a, b, out);
    case 2: return dispatch_field<Curve_BLS12381>(which, op, a, b, out);
  }
  return -1;
}

int hc_msm(int curve, int group, uint64_t srs_n, const uint8_t* pts, const uint8_t* scalars, uint64_t count,
           uint64_t batch, uint64_t base_offset, int c, int chunk, int bpc, int precomp, uint8_t* out, int* is_inf) {
  switch (curve) {
    case 0: return dispatch_msm<Curve_BN158>(group, srs_n, pts, scalars, count, batch, base_offset, c, chunk, bpc, precomp, out, is_inf);
    case 1: return dispatch_msm<Curve_BN254>(group, srs_n, pts, scalars, count, batch, base_offset, c, chunk, bpc, precomp, out, is_inf);
    case 2: return dispatch_msm<Curve_BLS12381>(group, srs_n, pts, scalars, count, batch, base_offset, c, chunk, bpc, precomp, out, is_inf);
  }
  return -1;
}

// G1 MSM through the direct-lookup table; wpt: windows per slot; spt: scalars per slot (used when wpt covers all windows)
int hc_msm_direct(int curve, uint64_t srs_n, const uint8_t* pts, const uint8_t* scalars, uint64_t count, uint64_t batch,
                  uint64_t base_offset, int c, int wpt, int spt, uint8_t* out, int* is_inf) {
  if (c < 2 || c > 12 || wpt < 1 || spt < 1) return -1;
  switch (curve) {
    case 0: return msm_direct_host<Curve_BN158>(srs_n, pts, scalars, count, batch, base_offset, c, wpt, spt, out, is_inf);
    case 1: return msm_direct_host<Curve_BN254>(srs_n, pts, scalars, count, batch, base_offset, c, wpt, spt, out, is_inf);
    case 2: return msm_direct_host<Curve_BLS12381>(srs_n, pts, scalars, count, batch, base_offset, c, wpt, spt, out, is_inf);
  }
  return -1;
}

}  // extern "C"
