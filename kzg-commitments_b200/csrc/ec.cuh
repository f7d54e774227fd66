// Short-Weierstrass group law, a = 0, generic over the coordinate field (Fp for G1, Fp2 for G2).
//
// Replaces, inside the MSM / fixed-base kernels, MIRACL's complete projective formulas
//   ECP_add  miracl-core/cpp/ecp.cpp:750-812     ECP_dbl  ecp.cpp:550-588
//   ECP2_add miracl-core/cpp/ecp2.cpp:413-502    ECP2_dbl ecp2.cpp:358-409
// with extended-Jacobian "XYZZ" coordinates (x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2): mixed add 8M+2S,
// full add 12M+2S, double 6M+3S.  These formulas are NOT complete, so every exceptional input the
// reference's complete formulas would absorb is branched on explicitly and yields the same group
// element: infinity operands, P + P (-> doubling) and P + (-P) (-> infinity).  The curve constant b
// never enters (a = 0), so the same code serves the D-type and M-type twists.
//
// Conventions: affine infinity is (0, 0) (not on any of the curves, b != 0); XYZZ infinity is ZZ = 0.
#pragma once
#include "field.cuh"
#include "fp2.cuh"

namespace kzgb200 {

template <class F>
struct Affine {
  F x, y;
  KZG_HD bool is_inf() const { return x.is_zero() && y.is_zero(); }
  KZG_HD static Affine inf() { return Affine{F::zero(), F::zero()}; }
};

template <class F>
struct XYZZ {
  F X, Y, ZZ, ZZZ;

  KZG_HD bool is_inf() const { return ZZ.is_zero(); }
  KZG_HD static XYZZ inf() { return XYZZ{F::zero(), F::zero(), F::zero(), F::zero()}; }
  KZG_HD static XYZZ from_affine(const Affine<F>& p) {
    if (p.is_inf()) return inf();
    return XYZZ{p.x, p.y, F::one(), F::one()};
  }

  // 2 * (affine p) -> XYZZ   (mdbl-2008-s-1)
  KZG_HDN static XYZZ dbl_affine(const F& x, const F& y) {
    F U = F::dbl(y);
    F V = F::sqr(U);
    F W = F::mul(U, V);
    F S = F::mul(x, V);
    F xx = F::sqr(x);
    F M = F::add(F::dbl(xx), xx);
    XYZZ r;
    r.X = F::sub(F::sqr(M), F::dbl(S));
    r.Y = F::sub(F::mul(M, F::sub(S, r.X)), F::mul(W, y));
    r.ZZ = V;
    r.ZZZ = W;
    return r;  // y == 0 gives ZZ = 0 = infinity, as it must
  }

  // this = 2 * this   (dbl-2008-s-1, a = 0)
  KZG_HDN void dbl() {
    if (is_inf()) return;
    F U = F::dbl(Y);
    F V = F::sqr(U);
    F W = F::mul(U, V);
    F S = F::mul(X, V);
    F xx = F::sqr(X);
    F M = F::add(F::dbl(xx), xx);
    F X3 = F::sub(F::sqr(M), F::dbl(S));
    F Y3 = F::sub(F::mul(M, F::sub(S, X3)), F::mul(W, Y));
    X = X3;
    Y = Y3;
    ZZ = F::mul(V, ZZ);
    ZZZ = F::mul(W, ZZZ);
  }

  // this += (x, y) with y already carrying the digit's sign   (madd-2008-s)
  KZG_HDN void madd(const F& x, const F& y) {
    if (x.is_zero() && y.is_zero()) return;  // affine infinity
    if (is_inf()) {
      X = x; Y = y; ZZ = F::one(); ZZZ = F::one();
      return;
    }
    F U2 = F::mul(x, ZZ);
    F S2 = F::mul(y, ZZZ);
    F Pd = F::sub(U2, X);
    F R = F::sub(S2, Y);
    if (Pd.is_zero()) {
      if (R.is_zero()) *this = dbl_affine(x, y);
      else *this = inf();
      return;
    }
    F PP = F::sqr(Pd);
    F PPP = F::mul(Pd, PP);
    F Q = F::mul(X, PP);
    F X3 = F::sub(F::sub(F::sqr(R), PPP), F::dbl(Q));
    F Y3 = F::sub(F::mul(R, F::sub(Q, X3)), F::mul(Y, PPP));
    X = X3;
    Y = Y3;
    ZZ = F::mul(ZZ, PP);
    ZZZ = F::mul(ZZZ, PPP);
  }
  KZG_HD void madd(const Affine<F>& p, bool negate) { madd(p.x, negate ? F::neg(p.y) : p.y); }

  // this += o   (add-2008-s)
  KZG_HDN void add(const XYZZ& o) {
    if (o.is_inf()) return;
    if (is_inf()) { *this = o; return; }
    F U1 = F::mul(X, o.ZZ);
    F U2 = F::mul(o.X, ZZ);
    F S1 = F::mul(Y, o.ZZZ);
    F S2 = F::mul(o.Y, ZZZ);
    F Pd = F::sub(U2, U1);
    F R = F::sub(S2, S1);
    if (Pd.is_zero()) {
      if (R.is_zero()) dbl();
      else *this = inf();
      return;
    }
    F PP = F::sqr(Pd);
    F PPP = F::mul(Pd, PP);
    F Q = F::mul(U1, PP);
    F X3 = F::sub(F::sub(F::sqr(R), PPP), F::dbl(Q));
    F Y3 = F::sub(F::mul(R, F::sub(Q, X3)), F::mul(S1, PPP));
    X = X3;
    Y = Y3;
    ZZ = F::mul(F::mul(ZZ, o.ZZ), PP);
    ZZZ = F::mul(F::mul(ZZZ, o.ZZZ), PPP);
  }

  KZG_HD void negate() { Y = F::neg(Y); }

  // x = X/ZZ, y = Y/ZZZ with one field inversion (replaces ECP_affine, ecp.cpp:329-348)
  KZG_HDN Affine<F> to_affine() const {
    if (is_inf()) return Affine<F>::inf();
    F i3 = F::inv(ZZZ);               // 1/Z^3
    F i1 = F::mul(ZZ, i3);            // 1/Z
    F i2 = F::sqr(i1);                // 1/Z^2
    return Affine<F>{F::mul(X, i2), F::mul(Y, i3)};
  }
};

// k * P by double-and-add over the bits of a little-endian word array (setup / fix-ups / tests)
template <class F>
KZG_HDN XYZZ<F> scalar_mul_bits(const Affine<F>& p, const uint32_t* k, int nbits) {
  XYZZ<F> acc = XYZZ<F>::inf();
  for (int i = nbits - 1; i >= 0; i--) {
    acc.dbl();
    if ((k[i >> 5] >> (i & 31)) & 1) acc.madd(p.x, p.y);
  }
  return acc;
}

}  // namespace kzgb200
