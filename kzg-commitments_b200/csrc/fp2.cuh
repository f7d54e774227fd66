// Fp2 = Fp[i]/(i^2 + 1) for the G2 path (replaces MIRACL FP2_mul/FP2_sqr/FP2_inv, miracl-core/cpp/fp2.cpp).
// Element a + b*i stored as (a, b); all three curve configs use QNRI = 0, i.e. i^2 = -1.
#pragma once
#include "field.cuh"

namespace kzgb200 {

template <class P>
struct Field2 {
  using Base = Field<P>;
  using Params = P;
  Base a, b;

  KZG_HD static Field2 zero() { return Field2{Base::zero(), Base::zero()}; }
  KZG_HD static Field2 one() { return Field2{Base::one(), Base::zero()}; }
  KZG_HD bool is_zero() const { return a.is_zero() && b.is_zero(); }
  KZG_HD bool operator==(const Field2& o) const { return a == o.a && b == o.b; }
  KZG_HD bool operator!=(const Field2& o) const { return !(*this == o); }

  KZG_HD static Field2 add(const Field2& x, const Field2& y) { return Field2{Base::add(x.a, y.a), Base::add(x.b, y.b)}; }
  KZG_HD static Field2 sub(const Field2& x, const Field2& y) { return Field2{Base::sub(x.a, y.a), Base::sub(x.b, y.b)}; }
  KZG_HD static Field2 dbl(const Field2& x) { return Field2{Base::dbl(x.a), Base::dbl(x.b)}; }
  KZG_HD static Field2 neg(const Field2& x) { return Field2{Base::neg(x.a), Base::neg(x.b)}; }

  // Karatsuba: 3 base multiplications
  KZG_HD static Field2 mul(const Field2& x, const Field2& y) {
    Base t0 = Base::mul(x.a, y.a);
    Base t1 = Base::mul(x.b, y.b);
    Base t2 = Base::mul(Base::add(x.a, x.b), Base::add(y.a, y.b));
    return Field2{Base::sub(t0, t1), Base::sub(Base::sub(t2, t0), t1)};
  }
  // (a+b)(a-b) + 2ab i : 2 base multiplications
  KZG_HD static Field2 sqr(const Field2& x) {
    Base t0 = Base::mul(Base::add(x.a, x.b), Base::sub(x.a, x.b));
    Base t1 = Base::mul(x.a, x.b);
    return Field2{t0, Base::dbl(t1)};
  }
  KZG_HDN static Field2 inv(const Field2& x) {
    Base n = Base::inv(Base::add(Base::sqr(x.a), Base::sqr(x.b)));
    return Field2{Base::mul(x.a, n), Base::neg(Base::mul(x.b, n))};
  }
};

}  // namespace kzgb200
