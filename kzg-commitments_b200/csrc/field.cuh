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

// 1 = mul() runs the Karatsuba / separated-reduction product where L % 4 == 0.  Measured on B200 (profiles/r1j):
// 12 % fewer wide multiplications but the same 30.9 G modmul/s and the same msm_accumulate time as the interleaved
// product, at 20 more registers — so the interleaved form stays the default and this one stays as a checked variant.
#ifndef KZG_KARATSUBA
#define KZG_KARATSUBA 0
#endif

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
  KZG_HDN static Field mul(Field a, Field b) {
#if KZG_KARATSUBA
    if constexpr (L % 4 == 0) return mul_karatsuba_inline(a, b);
    else return mul_inline(a, b);
#else
    return mul_inline(a, b);
#endif
  }
  // the body mul() uses, for call sites that want it fused (the modmul microbenchmark)
  KZG_HD static Field mul_best_inline(const Field& a, const Field& b) {
#if KZG_KARATSUBA
    if constexpr (L % 4 == 0) return mul_karatsuba_inline(a, b);
    else return mul_inline(a, b);
#else
    return mul_inline(a, b);
#endif
  }
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
  // ---- Karatsuba / separated-operand-scanning form of the same product (L divisible by 4).
  // The heavy integer-multiply pipe is the bound of every hot kernel (ncu: 93.6 % active in msm_accumulate,
  // profiles/r1i), while two thirds of the issue slots are idle.  One level of Karatsuba on the L x L
  // product trades L^2/4 wide multiplications (of 2L^2 + L) for ~6L additions, which run on the other
  // integer pipe:  a*b = z0 + ((al+ah)(bl+bh) - z0 - z2) 2^(16L) + z2 2^(32L),  three H x H products, H = L/2.
  // The Montgomery reduction then works on the low L words alone (L rows of m*p with two carry chains per
  // row on absolute-position accumulators) and the high L words are added at the end:
  //   (t + M p)/R = t_high + (t_low + M p)/R,   t_high < p,  (t_low + M p)/R <= p.
  static constexpr int H = L / 2;
  // z[0..2H) = x[0..H) * y[0..H); even and odd columns of x accumulate into two absolute-position arrays so every
  // row is two independent carry chains of fused multiply-adds
  KZG_HD static void mul_half(const uint32_t* x, const uint32_t* y, uint32_t* z) {
    uint32_t E[2 * H + 2], O[2 * H + 2];
#pragma unroll
    for (int k = 0; k < 2 * H + 2; k++) { E[k] = 0; O[k] = 0; }
#pragma unroll
    for (int j = 0; j < H; j += 2) {  // row 0
      E[j] = ptx::mul_lo(x[j], y[0]);
      E[j + 1] = ptx::mul_hi(x[j], y[0]);
      O[j + 1] = ptx::mul_lo(x[j + 1], y[0]);
      O[j + 2] = ptx::mul_hi(x[j + 1], y[0]);
    }
#pragma unroll
    for (int i = 1; i < H; i++) {
      // a product landing on words (q, q + 1) goes to E when q is even and to O when q is odd: every accumulator
      // word pair then stays one aligned 64-bit register pair for the fused wide multiply-add in all rows
      uint32_t* A = (i & 1) ? O : E;  // even columns of x: q = i + j has the parity of i
      uint32_t* B = (i & 1) ? E : O;  // odd columns
      const uint32_t yi = y[i];
      A[i] = ptx::mad_lo_cc(x[0], yi, A[i]);
      A[i + 1] = ptx::madc_hi_cc(x[0], yi, A[i + 1]);
#pragma unroll
      for (int j = 2; j < H; j += 2) {
        A[i + j] = ptx::madc_lo_cc(x[j], yi, A[i + j]);
        A[i + j + 1] = ptx::madc_hi_cc(x[j], yi, A[i + j + 1]);
      }
      A[i + H] = ptx::addc(A[i + H], 0);  // holds at most the previous row's carry
      B[i + 1] = ptx::mad_lo_cc(x[1], yi, B[i + 1]);
      B[i + 2] = ptx::madc_hi_cc(x[1], yi, B[i + 2]);
#pragma unroll
      for (int j = 3; j < H; j += 2) {
        B[i + j] = ptx::madc_lo_cc(x[j], yi, B[i + j]);
        B[i + j + 1] = ptx::madc_hi_cc(x[j], yi, B[i + j + 1]);
      }
      B[i + H + 1] = ptx::addc(B[i + H + 1], 0);
    }
    z[0] = E[0];
    z[1] = ptx::add_cc(E[1], O[1]);
#pragma unroll
    for (int k = 2; k < 2 * H - 1; k++) z[k] = ptx::addc_cc(E[k], O[k]);
    z[2 * H - 1] = ptx::addc(E[2 * H - 1], O[2 * H - 1]);
  }

  // t[0..2L) = a * b
  KZG_HD static void mul_wide_karatsuba(const Field& a, const Field& b, uint32_t* t) {
    uint32_t sa[H], sb[H], mid[2 * H + 1];
    mul_half(a.v, b.v, t);               // z0
    mul_half(a.v + H, b.v + H, t + L);   // z2
    sa[0] = ptx::add_cc(a.v[0], a.v[H]);
#pragma unroll
    for (int k = 1; k < H; k++) sa[k] = ptx::addc_cc(a.v[k], a.v[H + k]);
    const uint32_t ca = ptx::addc(0, 0);
    sb[0] = ptx::add_cc(b.v[0], b.v[H]);
#pragma unroll
    for (int k = 1; k < H; k++) sb[k] = ptx::addc_cc(b.v[k], b.v[H + k]);
    const uint32_t cb = ptx::addc(0, 0);
    mul_half(sa, sb, mid);               // (sa mod 2^(32H)) (sb mod 2^(32H))
    const uint32_t ma = 0u - ca, mb = 0u - cb;
    // + ca * sb * 2^(32H) + cb * sa * 2^(32H) + ca cb 2^(64H)
    mid[H] = ptx::add_cc(mid[H], sb[0] & ma);
#pragma unroll
    for (int k = 1; k < H; k++) mid[H + k] = ptx::addc_cc(mid[H + k], sb[k] & ma);
    mid[2 * H] = ptx::addc(ca & cb, 0);
    mid[H] = ptx::add_cc(mid[H], sa[0] & mb);
#pragma unroll
    for (int k = 1; k < H; k++) mid[H + k] = ptx::addc_cc(mid[H + k], sa[k] & mb);
    mid[2 * H] = ptx::addc(mid[2 * H], 0);
    // - z0 - z2   (the result a_l b_h + a_h b_l is non-negative and below 2^(64H + 1))
    mid[0] = ptx::sub_cc(mid[0], t[0]);
#pragma unroll
    for (int k = 1; k < 2 * H; k++) mid[k] = ptx::subc_cc(mid[k], t[k]);
    mid[2 * H] = ptx::subc(mid[2 * H], 0);
    mid[0] = ptx::sub_cc(mid[0], t[L]);
#pragma unroll
    for (int k = 1; k < 2 * H; k++) mid[k] = ptx::subc_cc(mid[k], t[L + k]);
    mid[2 * H] = ptx::subc(mid[2 * H], 0);
    // t += mid * 2^(32H)
    t[H] = ptx::add_cc(t[H], mid[0]);
#pragma unroll
    for (int k = 1; k <= 2 * H; k++) t[H + k] = ptx::addc_cc(t[H + k], mid[k]);
#pragma unroll
    for (int k = 3 * H + 1; k < 2 * L - 1; k++) t[k] = ptx::addc_cc(t[k], 0);
    t[2 * L - 1] = ptx::addc(t[2 * L - 1], 0);
  }

  // Montgomery reduction of t[0..2L) (t < p^2... any t < p * 2^(32L)) to the canonical residue t / R mod p
  KZG_HD static Field mont_reduce_wide(const uint32_t* t) {
    uint32_t E[2 * L + 2], O[2 * L + 2];
#pragma unroll
    for (int k = 0; k < 2 * L + 2; k++) { E[k] = k < L ? t[k] : 0u; O[k] = 0; }
#pragma unroll
    for (int i = 0; i < L; i++) {
      // complete word at position i; its carry enters the odd-column chain at position i + 1.  Accumulator choice
      // by the parity of the product's low word position, as in mul_half (keeps 64-bit register pairs aligned).
      uint32_t* A = (i & 1) ? O : E;
      uint32_t* B = (i & 1) ? E : O;
      A[i] = ptx::add_cc(A[i], B[i]);
      const uint32_t m = ptx::mul_lo(A[i], P::INV);
      B[i + 1] = ptx::madc_lo_cc(P::mod(1), m, B[i + 1]);
      B[i + 2] = ptx::madc_hi_cc(P::mod(1), m, B[i + 2]);
#pragma unroll
      for (int j = 3; j < L; j += 2) {
        B[i + j] = ptx::madc_lo_cc(P::mod(j), m, B[i + j]);
        B[i + j + 1] = ptx::madc_hi_cc(P::mod(j), m, B[i + j + 1]);
      }
      B[i + L + 1] = ptx::addc(B[i + L + 1], 0);
      A[i] = ptx::mad_lo_cc(P::mod(0), m, A[i]);  // == 0 by the choice of m; only its carry is wanted
      A[i + 1] = ptx::madc_hi_cc(P::mod(0), m, A[i + 1]);
#pragma unroll
      for (int j = 2; j < L; j += 2) {
        A[i + j] = ptx::madc_lo_cc(P::mod(j), m, A[i + j]);
        A[i + j + 1] = ptx::madc_hi_cc(P::mod(j), m, A[i + j + 1]);
      }
      A[i + L] = ptx::addc(A[i + L], 0);
    }
    Field r;
    r.v[0] = ptx::add_cc(E[L], O[L]);
#pragma unroll
    for (int k = 1; k < L - 1; k++) r.v[k] = ptx::addc_cc(E[L + k], O[L + k]);
    r.v[L - 1] = ptx::addc(E[2 * L - 1], O[2 * L - 1]);
    r.v[0] = ptx::add_cc(r.v[0], t[L]);
#pragma unroll
    for (int k = 1; k < L - 1; k++) r.v[k] = ptx::addc_cc(r.v[k], t[L + k]);
    r.v[L - 1] = ptx::addc(r.v[L - 1], t[2 * L - 1]);
    cond_sub_p(r);
    return r;
  }

  KZG_HD static Field mul_karatsuba_inline(const Field& a, const Field& b) {
    uint32_t t[2 * L];
    mul_wide_karatsuba(a, b, t);
    return mont_reduce_wide(t);
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

  // ---- Inverse by BATCHED DIVSTEPS ("safegcd", Bernstein-Yang 2019, in the 30-bit-limb variable-time form that
  // libsecp256k1's modinv32 popularised): f = p, g = a; 30 division steps at a time are run on the low words only and
  // collected in a 2x2 integer matrix t, which is then applied to the full-length (f, g) — exact division by 2^30 —
  // and to (d, e) modulo p, with d a = f, e a = g (mod p) throughout.  When g reaches 0, f = +-1 and a^-1 = +-d.
  // About 27 batches of ~560 native 32-bit instructions for BLS12-381 Fp (12 for BN158) against ~760 iterations of
  // multi-limb shifts / adds / compares for the binary GCD it replaces (kept below as inv_bingcd, the cross-check):
  // the single inversion at the end of every MSM was 90 us = 17 % of a degree-4096 commitment (profiles/r2_tail.md),
  // and the same routine sits in the setup normalisation and the window-table builds.
  // Variable time (the inputs are public group elements).  inv(0) = 0.  Limbs are 30-bit, the top one signed.
  static constexpr int N30 = (P::BITS + 2 + 29) / 30;
  KZG_HD static constexpr uint32_t mod_limb30(int i) {  // bits [30 i, 30 i + 30) of the modulus
    return (uint32_t)((((30 * i) / 32 < L ? (uint64_t)P::mod((30 * i) / 32 < L ? (30 * i) / 32 : 0) : 0ull) |
                       ((30 * i) / 32 + 1 < L ? (uint64_t)P::mod((30 * i) / 32 + 1 < L ? (30 * i) / 32 + 1 : 0) << 32 : 0ull)) >>
                      ((30 * i) % 32)) & 0x3fffffffu;
  }
  KZG_HD static int ctz_u32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
  }
  KZG_HDN static Field inv(const Field& a) {
    if (a.is_zero()) return zero();
    constexpr int N = N30;
    constexpr int32_t M30 = 0x3fffffff;
    constexpr uint32_t minv30 = (0u - P::INV) & 0x3fffffffu;  // p^-1 mod 2^30  (P::INV = -p^-1 mod 2^32)
    constexpr int MAX_BATCHES = (49 * P::BITS + 80) / 17 / 30 + 2;  // proven bound on the number of division steps
    int32_t f[N], g[N], d[N], e[N];
#pragma unroll
    for (int i = 0; i < N; i++) {
      const int w = (30 * i) >> 5, sh = (30 * i) & 31;
      uint64_t x = w < L ? (uint64_t)a.v[w < L ? w : 0] : 0ull;
      if (w + 1 < L) x |= (uint64_t)a.v[w + 1 < L ? w + 1 : 0] << 32;
      g[i] = (int32_t)((uint32_t)(x >> sh) & 0x3fffffffu);
      f[i] = (int32_t)mod_limb30(i);
      d[i] = 0;
      e[i] = 0;
    }
    e[0] = 1;
    int32_t eta = -1;
    for (int batch = 0; batch < MAX_BATCHES; batch++) {
      // 30 division steps on the low 32 bits of f and g; t = [u v; q r]
      uint32_t u = 1, v = 0, q = 0, r = 1;
      uint32_t fl = (uint32_t)f[0] | ((uint32_t)f[1] << 30), gl = (uint32_t)g[0] | ((uint32_t)g[1] << 30);
      int i = 30;
      for (;;) {
        const int zeros = ctz_u32(gl | (0xffffffffu << i));  // a sentinel bit stops the count at i
        gl >>= zeros; u <<= zeros; v <<= zeros; eta -= zeros; i -= zeros;
        if (i == 0) break;
        if (eta < 0) {
          eta = -eta;
          uint32_t t = fl; fl = gl; gl = 0u - t;
          t = u; u = q; q = 0u - t;
          t = v; v = r; r = 0u - t;
        }
        // f and g are odd: add the multiple w of f that clears the low min(eta + 1, i, 4) bits of g
        const int limit = (eta + 1) > i ? i : (eta + 1);
        const uint32_t m = (0xffffffffu >> (32 - limit)) & 15u;
        uint32_t w = fl + (((fl + 1u) & 4u) << 1);  // f^-1 mod 16
        w = (0u - w * gl) & m;
        gl += fl * w; q += u * w; r += v * w;
      }
      const int32_t tu = (int32_t)u, tv = (int32_t)v, tq = (int32_t)q, tr = (int32_t)r;
      // (d, e) <- t (d, e) / 2^30 mod p: a multiple of p is added so that the division is exact
      {
        const int32_t sd = d[N - 1] >> 31, se = e[N - 1] >> 31;
        int32_t md = (tu & sd) + (tv & se), me = (tq & sd) + (tr & se);
        int64_t cd = (int64_t)tu * d[0] + (int64_t)tv * e[0], ce = (int64_t)tq * d[0] + (int64_t)tr * e[0];
        md -= (int32_t)((minv30 * (uint32_t)cd + (uint32_t)md) & (uint32_t)M30);
        me -= (int32_t)((minv30 * (uint32_t)ce + (uint32_t)me) & (uint32_t)M30);
        cd += (int64_t)(int32_t)mod_limb30(0) * md;
        ce += (int64_t)(int32_t)mod_limb30(0) * me;
        cd >>= 30;
        ce >>= 30;
#pragma unroll
        for (int k = 1; k < N; k++) {
          cd += (int64_t)tu * d[k] + (int64_t)tv * e[k] + (int64_t)(int32_t)mod_limb30(k) * md;
          ce += (int64_t)tq * d[k] + (int64_t)tr * e[k] + (int64_t)(int32_t)mod_limb30(k) * me;
          d[k - 1] = (int32_t)cd & M30; cd >>= 30;
          e[k - 1] = (int32_t)ce & M30; ce >>= 30;
        }
        d[N - 1] = (int32_t)cd;
        e[N - 1] = (int32_t)ce;
      }
      // (f, g) <- t (f, g) / 2^30, exact
      {
        int64_t cf = (int64_t)tu * f[0] + (int64_t)tv * g[0], cg = (int64_t)tq * f[0] + (int64_t)tr * g[0];
        cf >>= 30;
        cg >>= 30;
        int32_t nz = 0;
#pragma unroll
        for (int k = 1; k < N; k++) {
          cf += (int64_t)tu * f[k] + (int64_t)tv * g[k];
          cg += (int64_t)tq * f[k] + (int64_t)tr * g[k];
          f[k - 1] = (int32_t)cf & M30; cf >>= 30;
          g[k - 1] = (int32_t)cg & M30; cg >>= 30;
          nz |= g[k - 1];
        }
        f[N - 1] = (int32_t)cf;
        g[N - 1] = (int32_t)cg;
        nz |= g[N - 1];
        if (nz == 0) break;
      }
    }
    // f = +-1 and d a = f (mod p), d in (-2p, p): a^-1 = sign(f) d, brought into [0, p)
    if (f[N - 1] < 0) {
      int32_t c = 0;
#pragma unroll
      for (int k = 0; k < N - 1; k++) { const int32_t t = c - d[k]; d[k] = t & M30; c = t >> 30; }
      d[N - 1] = c - d[N - 1];
    }
    for (int pass = 0; pass < 2; pass++) {  // + p while negative (at most twice)
      if (d[N - 1] < 0) {
        int32_t c = 0;
#pragma unroll
        for (int k = 0; k < N - 1; k++) { const int32_t t = d[k] + (int32_t)mod_limb30(k) + c; d[k] = t & M30; c = t >> 30; }
        d[N - 1] += (int32_t)mod_limb30(N - 1) + c;
      }
    }
    {  // - p if that stays non-negative
      int32_t t[N], c = 0;
#pragma unroll
      for (int k = 0; k < N - 1; k++) { const int32_t x = d[k] - (int32_t)mod_limb30(k) + c; t[k] = x & M30; c = x >> 30; }
      t[N - 1] = d[N - 1] - (int32_t)mod_limb30(N - 1) + c;
      if (t[N - 1] >= 0) {
#pragma unroll
        for (int k = 0; k < N; k++) d[k] = t[k];
      }
    }
    Field res;
#pragma unroll
    for (int j = 0; j < L; j++) {  // 30-bit limbs -> 32-bit words
      const int k = (32 * j) / 30, sh = (32 * j) % 30;
      uint32_t x = k < N ? (uint32_t)d[k < N ? k : 0] >> sh : 0u;
      if (k + 1 < N) x |= (uint32_t)d[k + 1 < N ? k + 1 : 0] << (30 - sh);
      if (k + 2 < N && 60 - sh < 32) x |= (uint32_t)d[k + 2 < N ? k + 2 : 0] << (60 - sh);
      res.v[j] = x;
    }
    // res = (aR)^-1 as a plain residue; two products by R^2 give a^-1 * R
    res = mul(res, r2());
    return mul(res, r2());
  }

  // The binary extended Euclidean algorithm on the raw limbs (shifts, adds, compares; ~2*BITS short iterations): the
  // inversion of round 1, kept as an independent formulation for the self-tests.
  KZG_HDN static Field inv_bingcd(const Field& a) {
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
