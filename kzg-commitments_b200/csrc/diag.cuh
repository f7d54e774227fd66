// Diagnostics that ship with the library: integer-pipe microbenchmarks (the roofline denominator of
// SURVEY.md 8(d) must be MEASURED on the box) and an on-device self-test of the field / group-law
// layers against independent formulations.
#pragma once
#include "ec.cuh"

namespace kzgb200 {

// ---------------------------------------------------------------- microbenchmarks
#if defined(__CUDACC__)
// kind 0: rows of 8 lo/hi multiply-add pairs with one carry chain per row (what mul() issues);
// 4 independent rows per thread.  products = threads * iters * 32.
static __global__ void __launch_bounds__(256) bench_wide_chain(int iters, uint32_t seed, uint32_t* sink) {
  uint32_t acc[4][16];
  uint32_t a[8];
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
  for (int i = 0; i < 8; i++) a[i] = seed * (tid + i + 1) + 0x9e3779b9u * i;
#pragma unroll
  for (int r = 0; r < 4; r++)
#pragma unroll
    for (int i = 0; i < 16; i++) acc[r][i] = tid ^ (r * 16 + i);
  uint32_t b = seed ^ tid;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int r = 0; r < 4; r++) {
      acc[r][0] = ptx::mad_lo_cc(a[0], b, acc[r][0]);
      acc[r][1] = ptx::madc_hi_cc(a[0], b, acc[r][1]);
#pragma unroll
      for (int j = 1; j < 8; j++) {
        acc[r][2 * j] = ptx::madc_lo_cc(a[j], b, acc[r][2 * j]);
        acc[r][2 * j + 1] = ptx::madc_hi_cc(a[j], b, acc[r][2 * j + 1]);
      }
    }
    b = b * 1664525u + 1013904223u;
  }
  uint32_t x = 0;
#pragma unroll
  for (int r = 0; r < 4; r++)
#pragma unroll
    for (int i = 0; i < 16; i++) x ^= acc[r][i];
  if (x == 0x12345678u) sink[0] = x;
}

// kind 1: independent 32-bit IMAD (mad.lo.u32), 16 accumulators per thread. ops = threads*iters*16
static __global__ void __launch_bounds__(256) bench_imad32(int iters, uint32_t seed, uint32_t* sink) {
  uint32_t acc[16];
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] = tid + i;
  uint32_t b = seed ^ tid;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = acc[i] * b + (uint32_t)i;
    b += 2;
  }
  uint32_t x = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) x ^= acc[i];
  if (x == 0x12345678u) sink[0] = x;
}

// kind 2/3: Montgomery products, two independent dependency chains per thread. ops = threads*iters*2
template <class F>
__global__ void __launch_bounds__(256) bench_modmul(int iters, uint32_t seed, uint32_t* sink) {
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  F x = F::one(), y = F::r2(), z = F::one();
  x.v[0] ^= (seed * (tid + 1)) & 0xffff;
  z.v[0] ^= (seed * (tid + 7)) & 0xffff;
  for (int it = 0; it < iters; it++) {
    x = F::mul_best_inline(x, y);
    z = F::mul_best_inline(z, y);
  }
  uint32_t o = 0;
#pragma unroll
  for (int i = 0; i < F::L; i++) o ^= x.v[i] ^ z.v[i];
  if (o == 0x12345678u) sink[0] = o;
}

#endif  // __CUDACC__

// ---------------------------------------------------------------- self-test
struct Lcg {
  uint32_t s;
  KZG_HD uint32_t next() { s = s * 1664525u + 1013904223u; return s ^ (s >> 15); }
};

template <class F>
KZG_HD F rand_fe(Lcg& g) {
  F a;
  for (int i = 0; i < F::L; i++) a.v[i] = g.next();
  return F::to_mont(a);  // canonical for any input
}

template <class F>
KZG_HDN int check_field(Lcg& g) {
  int bad = 0;
  F a = rand_fe<F>(g), b = rand_fe<F>(g), c = rand_fe<F>(g);
  F lhs = F::sqr(F::add(a, b));
  F rhs = F::add(F::add(F::sqr(a), F::dbl(F::mul(a, b))), F::sqr(b));
  bad += !(lhs == rhs);
  bad += !(F::mul(a, b) == F::mul(b, a));
  bad += !(F::mul(F::mul(a, b), c) == F::mul(a, F::mul(b, c)));
  bad += !(F::mul(a, F::sub(b, c)) == F::sub(F::mul(a, b), F::mul(a, c)));
  bad += !(F::add(F::sub(a, b), b) == a);
  bad += !(F::add(a, F::neg(a)).is_zero());
  bad += !(F::to_mont(F::from_mont(a)) == a);
  if (!a.is_zero()) bad += !(F::mul(a, F::inv(a)) == F::one());
  bad += !(F::inv(a) == F::inv_fermat(a));  // divstep (safegcd) inverse against the Fermat ladder
  bad += !(F::inv(a) == F::inv_bingcd(a));  // ... and against the binary extended GCD
  // the Karatsuba / separated-reduction product (field.cuh, KZG_KARATSUBA) against the interleaved one
  if constexpr (F::L % 4 == 0) {
    const F m1 = F::neg(F::one());
    bad += !(F::mul_karatsuba_inline(a, b) == F::mul_inline(a, b));
    bad += !(F::mul_karatsuba_inline(a, a) == F::mul_inline(a, a));
    bad += !(F::mul_karatsuba_inline(m1, m1) == F::mul_inline(m1, m1));
    bad += !(F::mul_karatsuba_inline(F::r2(), m1) == F::mul_inline(F::r2(), m1));
  }
  return bad;
}

template <class F>
KZG_HDN bool same_point(const XYZZ<F>& p, const XYZZ<F>& q) {
  if (p.is_inf() || q.is_inf()) return p.is_inf() && q.is_inf();
  return F::mul(p.X, q.ZZ) == F::mul(q.X, p.ZZ) && F::mul(p.Y, q.ZZZ) == F::mul(q.Y, p.ZZZ);
}

// F2 = field of coordinates, b = curve constant (Montgomery) for the on-curve check
template <class F>
KZG_HDN int check_group(const Affine<F>& gen, const F& bcoef, const uint32_t* order, int order_bits, Lcg& g) {
  int bad = 0;
  uint32_t k1[2] = {g.next(), g.next() & 0xffff}, k2[2] = {g.next(), g.next() & 0xffff};
  XYZZ<F> P = scalar_mul_bits<F>(gen, k1, 48), Q = scalar_mul_bits<F>(gen, k2, 48);
  Affine<F> Qa = Q.to_affine(), Pa = P.to_affine();
  // on curve
  bad += !(F::sqr(Pa.y) == F::add(F::mul(F::sqr(Pa.x), Pa.x), bcoef));
  // mixed add == full add
  XYZZ<F> s1 = P; s1.madd(Qa.x, Qa.y);
  XYZZ<F> s2 = P; s2.add(Q);
  bad += !same_point(s1, s2);
  // commutativity, P + Q - Q == P
  XYZZ<F> s3 = Q; s3.add(P);
  bad += !same_point(s1, s3);
  XYZZ<F> s4 = s1; s4.madd(Qa, true);
  bad += !same_point(s4, P);
  // doubling three ways
  XYZZ<F> d1 = P; d1.dbl();
  XYZZ<F> d2 = P; d2.add(P);
  XYZZ<F> d3 = P; d3.madd(Pa.x, Pa.y);
  bad += !same_point(d1, d2);
  bad += !same_point(d1, d3);
  bad += !same_point(d1, XYZZ<F>::dbl_affine(Pa.x, Pa.y));
  // P - P = inf, inf + P = P
  XYZZ<F> z = P; z.madd(Pa, true);
  bad += !z.is_inf();
  z.madd(Pa.x, Pa.y);
  bad += !same_point(z, P);
  XYZZ<F> z2 = P; XYZZ<F> nP = P; nP.negate(); z2.add(nP);
  bad += !z2.is_inf();
  // (k1 + k2) G == P + Q   (k1 + k2 < 2^49)
  uint64_t ks = ((uint64_t)k1[1] << 32 | k1[0]) + ((uint64_t)k2[1] << 32 | k2[0]);
  uint32_t k3[2] = {(uint32_t)ks, (uint32_t)(ks >> 32)};
  bad += !same_point(scalar_mul_bits<F>(gen, k3, 50), s1);
  // [r]G = inf
  bad += !scalar_mul_bits<F>(gen, order, order_bits).is_inf();
  return bad;
}

template <class C>
KZG_HD int selftest_one(uint32_t seed) {
  using Fp = Field<typename C::FpP>;
  using Fp2 = Field2<typename C::FpP>;
  using Fr = Field<typename C::FrP>;
  Lcg g{seed * 2654435761u + 12345u};
  int bad = 0;
  bad += check_field<Fp>(g);
  bad += check_field<Fr>(g);
  {
    Fp2 a{rand_fe<Fp>(g), rand_fe<Fp>(g)}, b{rand_fe<Fp>(g), rand_fe<Fp>(g)};
    bad += !(Fp2::sqr(a) == Fp2::mul(a, a));
    bad += !(Fp2::mul(a, b) == Fp2::mul(b, a));
    bad += !(Fp2::mul(a, Fp2::inv(a)) == Fp2::one());
    bad += !(Fp2::sqr(Fp2::add(a, b)) == Fp2::add(Fp2::add(Fp2::sqr(a), Fp2::dbl(Fp2::mul(a, b))), Fp2::sqr(b)));
  }
  uint32_t order[Fr::L];
  for (int i = 0; i < Fr::L; i++) order[i] = C::FrP::mod(i);
  {
    Affine<Fp> gen;
    Fp bc;
    for (int i = 0; i < Fp::L; i++) { gen.x.v[i] = C::g1x(i); gen.y.v[i] = C::g1y(i); bc.v[i] = C::b1(i); }
    bad += check_group<Fp>(gen, bc, order, C::SCALAR_BITS, g);
  }
  {
    Affine<Fp2> gen;
    Fp2 bc;
    for (int i = 0; i < Fp::L; i++) {
      gen.x.a.v[i] = C::g2xa(i); gen.x.b.v[i] = C::g2xb(i);
      gen.y.a.v[i] = C::g2ya(i); gen.y.b.v[i] = C::g2yb(i);
      bc.a.v[i] = C::b2a(i); bc.b.v[i] = C::b2b(i);
    }
    bad += check_group<Fp2>(gen, bc, order, C::SCALAR_BITS, g);
  }
  return bad;
}

#if defined(__CUDACC__)
template <class C>
__global__ void selftest_kernel(uint32_t seed, int* failures) {
  int bad = selftest_one<C>(seed + blockIdx.x * blockDim.x + threadIdx.x);
  if (bad) atomicAdd(failures, bad);
}
#endif

}  // namespace kzgb200
