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
  std::vector<uint32_t> counts(nbuckets + 1, 0), starts(nbuckets + 1), task_first(nbuckets + 1), cursor(nbuckets + 1, 0),
      sorted(records);
  for (uint64_t t = 0; t < count * batch; t++) msm_digits_thread<C>(sh, scalars, digits.data(), counts.data(), t);
  uint32_t a = 0, b = 0;
  for (uint64_t i = 0; i < nbuckets; i++) {
    starts[i] = a; task_first[i] = b;
    a += counts[i]; b += (counts[i] + chunk - 1) / chunk;
  }
  starts[nbuckets] = a; task_first[nbuckets] = b;
  for (uint64_t t = 0; t < records; t++) msm_scatter_thread(sh, digits.data(), starts.data(), cursor.data(), sorted.data(), t);
  std::vector<XYZZ<F>> partial(b + 1);
  for (uint32_t t = 0; t < b; t++)
    msm_accumulate_thread<F>(nbuckets, task_first.data(), starts.data(), sorted.data(),
                             precomp ? table.data() : bases.data(), partial.data(), t);
  if (bpc > sh.nb) bpc = sh.nb;
  const int cpg = sh.nb / bpc;
  const uint64_t groups = sh.total_groups();
  std::vector<XYZZ<F>> chunk_out(groups * cpg), group_out(groups);
  for (uint64_t t = 0; t < groups * cpg; t++)
    msm_reduce_chunk_thread<F>(sh.nb, bpc, task_first.data(), partial.data(), chunk_out.data(), t);
  for (uint64_t g = 0; g < groups; g++) {
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int j = 0; j < cpg; j++) acc.add(chunk_out[g * cpg + j]);
    group_out[g] = acc;
  }
  for (uint64_t m = 0; m < batch; m++)
    msm_finish_thread<F>(sh.groups_per_msm, sh.c, group_out.data(), out, is_inf, M, (XYZZ<F>*)nullptr, m);
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

}  // extern "C"
