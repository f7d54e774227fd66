// Prime-field arithmetic in Montgomery form on 32-bit saturated limbs (R = 2^(32*L), L even).
//
// Replaces, for the CUDA path, what the reference obtains from MIRACL's BIG/FP layer:
//   FP_mul / FP_sqr  (miracl-core/cpp/fp.cpp:396-415 -> BIG_mul + BIG_monty, big.cpp:570-680,782-897)
//   FP_add / FP_sub / FP_neg (fp.cpp:485-610),  FP_inv (fp.cpp:817-840)
// and NTL's ZZ_p mul/add for Fr.  MIRACL keeps 56/58-bit unsaturated limbs with lazy "excess"
// tracking on 64-bit CPUs; here every value is kept canonical in [0, p) on L 32-bit limbs so that
// equality tests (needed by the exceptional cases of the group law) are plain limb compares.
//
// mul(): operand-scanning Montgomery product with two accumulators ("even"/"odd" aligned) so that
// the lo/hi halves of each 32x32 product land on a 64-bit aligned register pair with one carry
// chain per row; ptxas turns each mad.lo.cc/madc.hi.cc pair into one IMAD.WIDE.U32(.X).
// Cost: 2*L^2 wide multiply-adds + L single multiplies (m = t0 * INV) per product.
#pragma once
#include "ptx.cuh"

namespace kzgb200 {

template <class P>
struct alignas((P::L % 4 == 0) ? 16 : 8) Field {  // aligned for 128-bit (64-bit for 6 limbs) loads and stores
  static constexpr int L = P::L;
  using Params = P;
  uint32_t v[L];

  KZG_HD static Field zero() {
    Field r;
#pragma unroll
    for (int i = 0; i < L; i++) r.v[i] = 0;
    return r;
  }
  KZG_HD static Field one() {
    Field r;
#pragma unroll
    for (int i = 0; i < L; i++) r.v[i] = P::one(i);
    return r;
  }
  KZG_HD static Field r2() {
    Field r;
#pragma unroll
    for (int i = 0; i < L; i++) r.v[i] = P::r2(i);
    return r;
  }
  KZG_HD bool is_zero() const {
    uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < L; i++) o |= v[i];
    return o == 0;
  }
  KZG_HD bool operator==(const Field& b) const {
    uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < L; i++) o |= v[i] ^ b.v[i];
    return o == 0;
  }
  KZG_HD bool operator!=(const Field& b) const { return !(*this == b); }

  // r = a - p if a >= p else a   (a < 2p, possibly with an extra carry bit `hi`)
  KZG_HD static void cond_sub_p(Field& a, uint32_t hi = 0) {
    uint32_t t[L];
    t[0] = ptx::sub_cc(a.v[0], P::mod(0));
#pragma unroll
    for (int i = 1; i < L; i++) t[i] = ptx::subc_cc(a.v[i], P::mod(i));
    uint32_t borrow = ptx::subc(hi, 0);  // 0 if (hi:a) >= p, 0xffffffff otherwise
    bool keep = borrow != 0;
#pragma unroll
    for (int i = 0; i < L; i++) a.v[i] = keep ? a.v[i] : t[i];
  }

  KZG_HD static Field add(const Field& a, const Field& b) {
    Field r;
    r.v[0] = ptx::add_cc(a.v[0], b.v[0]);
#pragma unroll
    for (int i = 1; i < L; i++) r.v[i] = ptx::addc_cc(a.v[i], b.v[i]);
    uint32_t hi = ptx::addc(0, 0);
    cond_sub_p(r, hi);
    return r;
  }
  KZG_HD static Field dbl(const Field& a) { return add(a, a); }

  KZG_HD static Field sub(const Field& a, const Field& b) {
    Field r;
    r.v[0] = ptx::sub_cc(a.v[0], b.v[0]);
#pragma unroll
    for (int i = 1; i < L; i++) r.v[i] = ptx::subc_cc(a.v[i], b.v[i]);
    uint32_t mask = ptx::subc(0, 0);  // all ones when a < b
    r.v[0] = ptx::add_cc(r.v[0], P::mod(0) & mask);
#pragma unroll
    for (int i = 1; i < L - 1; i++) r.v[i] = ptx::addc_cc(r.v[i], P::mod(i) & mask);
    r.v[L - 1] = ptx::addc(r.v[L - 1], P::mod(L - 1) & mask);
    return r;
  }
  KZG_HD static Field neg(const Field& a) { return sub(zero(), a); }

  // ---- Montgomery product a*b/R mod p, inputs and output canonical.
  // mul() is ONE out-of-line routine per field (operands by value: ptxas passes the 2*L words in
  // registers, no stack traffic).  Inlining the ~2L^2 multiply-adds at every call site made the
  // group-law code tens of KiB per routine and the bucket-accumulation kernel instruction-fetch
  // bound (ncu: stall_no_instruction 4.3 of 9.4 cycles per issue, profiles/r1_accumulate_v1.md);
  // with the call the hot loop fits the instruction cache.  mul_inline() is the same body for
  // call sites that want it fused (the modmul microbenchmark).
  KZG_HDN static Field mul(Field a, Field b) { return mul_inline(a, b); }
  KZG_HD static Field mul_inline(const Field& a, const Field& b) {
    static_assert(L % 2 == 0, "even limb count required");
    uint32_t ev[L], od[L];
    // row 0: ev = a_even * b0 (pairs j,j+1 for even j) ; od = a_odd * b0 (value at limb offset 1)
    {
      const uint32_t bi = b.v[0];
#pragma unroll
      for (int j = 0; j < L; j += 2) {
        ev[j] = ptx::mul_lo(a.v[j], bi);
        ev[j + 1] = ptx::mul_hi(a.v[j], bi);
        od[j] = ptx::mul_lo(a.v[j + 1], bi);
        od[j + 1] = ptx::mul_hi(a.v[j + 1], bi);
      }
      reduce_row(ev, od);
    }
#pragma unroll
    for (int i = 1; i < L; i++) {
      if (i & 1) mad_row(od, ev, a, b.v[i]);
      else mad_row(ev, od, a, b.v[i]);
    }
    // rows alternate; L even => the last row used lo = od, hi = ev.  result = od/2^32 + ev
    Field r;
    r.v[0] = ptx::add_cc(od[1], ev[0]);
#pragma unroll
    for (int k = 1; k < L - 1; k++) r.v[k] = ptx::addc_cc(od[k + 1], ev[k]);
    r.v[L - 1] = ptx::addc(ev[L - 1], 0);
    cond_sub_p(r);
    return r;
  }
  KZG_HD static Field sqr(const Field& a) { return mul(a, a); }

  // R^2 is the row operand so that ANY 32L-bit input (not only canonical ones) comes out canonical:
  // the running value stays below r2 + p, independent of the digits of `a`.
  KZG_HD static Field to_mont(const Field& a) { return mul(r2(), a); }
  KZG_HD static Field from_mont(const Field& a) {
    Field o = zero();
    o.v[0] = 1;
    return mul(a, o);
  }

  // Inverse by the binary extended Euclidean algorithm on the raw limbs (shifts, adds, compares: no
  // multiplications, ~2*BITS short iterations) followed by two Montgomery products to restore the
  // R factor.  A single thread's Fermat ladder is ~1.5*BITS dependent modmuls (0.65 ms for BLS12-381 Fp,
  // profiles/r1c); this is several times shorter, which matters because the MSM finish is ONE inversion
  // on ONE thread.  Variable time (the inputs here are public group elements).  inv(0) = 0.
  KZG_HDN static Field inv(const Field& a) {
    if (a.is_zero()) return zero();
    uint32_t u[L], w[L], x1[L], x2[L];
#pragma unroll
    for (int i = 0; i < L; i++) { u[i] = a.v[i]; w[i] = P::mod(i); x1[i] = 0; x2[i] = 0; }
    x1[0] = 1;
    // invariants: x1 * a == u, x2 * a == w (mod p), all values < p
    for (int guard = 0; guard < 4 * P::BITS + 8; guard++) {
      if (is_one(u) || is_one(w)) break;
      while ((u[0] & 1u) == 0) { shr1(u, 0); halve_mod(x1); }
      while ((w[0] & 1u) == 0) { shr1(w, 0); halve_mod(x2); }
      if (geq(u, w)) { sub_raw(u, w); sub_mod(x1, x2); }
      else { sub_raw(w, u); sub_mod(x2, x1); }
    }
    Field r;
    const bool take1 = is_one(u);
#pragma unroll
    for (int i = 0; i < L; i++) r.v[i] = take1 ? x1[i] : x2[i];
    // r = (aR)^-1 as a plain residue; two products by R^2 give a^-1 * R
    r = mul(r, r2());
    return mul(r, r2());
  }
  // a^(p-2) by square-and-multiply over the constant exponent (what FP_inv does, fp.cpp:817-840); kept as
  // the independent formulation the device self-test compares inv() against.
  KZG_HDN static Field inv_fermat(const Field& a) {
    Field r = one();
    for (int i = P::BITS - 1; i >= 0; i--) {
      r = sqr(r);
      uint32_t w = mod_minus_2_word(i >> 5);
      if ((w >> (i & 31)) & 1) r = mul(r, a);
    }
    return r;
  }

 private:
  // ---- raw multi-limb helpers for inv()
  KZG_HD static bool is_one(const uint32_t* a) {
    uint32_t o = a[0] ^ 1u;
#pragma unroll
    for (int i = 1; i < L; i++) o |= a[i];
    return o == 0;
  }
  KZG_HD static void shr1(uint32_t* a, uint32_t top) {  // a = (top:a) >> 1
#pragma unroll
    for (int i = 0; i < L - 1; i++) a[i] = (a[i] >> 1) | (a[i + 1] << 31);
    a[L - 1] = (a[L - 1] >> 1) | (top << 31);
  }
  KZG_HD static void halve_mod(uint32_t* x) {  // x = x/2 mod p
    uint32_t top = 0;
    if (x[0] & 1u) {
      x[0] = ptx::add_cc(x[0], P::mod(0));
#pragma unroll
      for (int i = 1; i < L; i++) x[i] = ptx::addc_cc(x[i], P::mod(i));
      top = ptx::addc(0, 0);
    }
    shr1(x, top);
  }
  KZG_HD static bool geq(const uint32_t* a, const uint32_t* b) {  // a >= b
    uint32_t t = ptx::sub_cc(a[0], b[0]);
#pragma unroll
    for (int i = 1; i < L; i++) t = ptx::subc_cc(a[i], b[i]);
    (void)t;
    return ptx::subc(0, 0) == 0;
  }
  KZG_HD static void sub_raw(uint32_t* a, const uint32_t* b) {  // a -= b, a >= b
    a[0] = ptx::sub_cc(a[0], b[0]);
#pragma unroll
    for (int i = 1; i < L - 1; i++) a[i] = ptx::subc_cc(a[i], b[i]);
    a[L - 1] = ptx::subc(a[L - 1], b[L - 1]);
  }
  KZG_HD static void sub_mod(uint32_t* a, const uint32_t* b) {  // a = a - b mod p
    a[0] = ptx::sub_cc(a[0], b[0]);
#pragma unroll
    for (int i = 1; i < L; i++) a[i] = ptx::subc_cc(a[i], b[i]);
    const uint32_t mask = ptx::subc(0, 0);
    a[0] = ptx::add_cc(a[0], P::mod(0) & mask);
#pragma unroll
    for (int i = 1; i < L - 1; i++) a[i] = ptx::addc_cc(a[i], P::mod(i) & mask);
    a[L - 1] = ptx::addc(a[L - 1], P::mod(L - 1) & mask);
  }
  KZG_HD static uint32_t mod_minus_2_word(int k) {
    uint32_t w = 0;
#pragma unroll
    for (int i = 0; i < L; i++) w = (i == k) ? P::mod_minus_2(i) : w;
    return w;
  }

  // lo: array aligned at limb 0 for this row, hi: array whose limb k sits at position k+1.
  // Adds m*p with m chosen so that lo[0] becomes 0; the carry out of lo goes to hi[L-1]
  // (same absolute position).
  KZG_HD static void reduce_row(uint32_t* lo, uint32_t* hi) {
    const uint32_t m = ptx::mul_lo(lo[0], P::INV);
    hi[0] = ptx::mad_lo_cc(P::mod(1), m, hi[0]);
    hi[1] = ptx::madc_hi_cc(P::mod(1), m, hi[1]);
#pragma unroll
    for (int j = 3; j < L; j += 2) {
      hi[j - 1] = ptx::madc_lo_cc(P::mod(j), m, hi[j - 1]);
      hi[j] = ptx::madc_hi_cc(P::mod(j), m, hi[j]);
    }
    lo[0] = ptx::mad_lo_cc(P::mod(0), m, lo[0]);
    lo[1] = ptx::madc_hi_cc(P::mod(0), m, lo[1]);
#pragma unroll
    for (int j = 2; j < L; j += 2) {
      lo[j] = ptx::madc_lo_cc(P::mod(j), m, lo[j]);
      lo[j + 1] = ptx::madc_hi_cc(P::mod(j), m, lo[j + 1]);
    }
    hi[L - 1] = ptx::addc(hi[L - 1], 0);
  }

  // One interior row.  On entry the running value is  T = hi_prev/2^32 + lo  where `lo` is the array
  // that was odd-aligned in the previous row and `hi` the one that was even-aligned (hi[0] == 0).
  // On exit T' = (T + a*bi + m*p)/2^32 = lo/2^32 + hi with lo[0] == 0.
  KZG_HD static void mad_row(uint32_t* lo, uint32_t* hi, const Field& a, uint32_t bi) {
    // fold hi[1] (position 0 after the shift) into lo[0]; its carry enters the odd chain below
    lo[0] = ptx::add_cc(lo[0], hi[1]);
    // odd columns: hi'[j-1],hi'[j] = a[j]*bi + hi[j+1],hi[j+2]   (shift by two limbs for free)
#pragma unroll
    for (int j = 1; j < L - 1; j += 2) {
      hi[j - 1] = ptx::madc_lo_cc(a.v[j], bi, hi[j + 1]);
      hi[j] = ptx::madc_hi_cc(a.v[j], bi, hi[j + 2]);
    }
    hi[L - 2] = ptx::madc_lo_cc(a.v[L - 1], bi, 0);
    hi[L - 1] = ptx::madc_hi(a.v[L - 1], bi, 0);
    // even columns
    lo[0] = ptx::mad_lo_cc(a.v[0], bi, lo[0]);
    lo[1] = ptx::madc_hi_cc(a.v[0], bi, lo[1]);
#pragma unroll
    for (int j = 2; j < L; j += 2) {
      lo[j] = ptx::madc_lo_cc(a.v[j], bi, lo[j]);
      lo[j + 1] = ptx::madc_hi_cc(a.v[j], bi, lo[j + 1]);
    }
    hi[L - 1] = ptx::addc(hi[L - 1], 0);
    reduce_row(lo, hi);
  }
};

}  // namespace kzgb200
