// dd.cuh — double-double arithmetic (error-free transformations) for the small, ill-conditioned parts of the
// regression: Gram matrix of the monomial basis, its Cholesky factor, R^-1 and the orthonormalised basis rows.
// Compiled with -fmad=false: no contraction may touch these sequences; FMAs are explicit.
#pragma once

namespace lsmc {

struct dd {
  double hi, lo;
};

__host__ __device__ __forceinline__ dd two_sum(double a, double b) {
  double s = a + b;
  double bb = s - a;
  double e = (a - (s - bb)) + (b - bb);
  return {s, e};
}
__host__ __device__ __forceinline__ dd quick_two_sum(double a, double b) {
  double s = a + b;
  return {s, b - (s - a)};
}
__host__ __device__ __forceinline__ dd two_prod(double a, double b) {
  double p = a * b;
#ifdef __CUDA_ARCH__
  double e = __fma_rn(a, b, -p);
#else
  double e = __builtin_fma(a, b, -p);
#endif
  return {p, e};
}
__host__ __device__ __forceinline__ dd dd_add(dd a, dd b) {
  dd s = two_sum(a.hi, b.hi);
  dd t = two_sum(a.lo, b.lo);
  s.lo += t.hi;
  s = quick_two_sum(s.hi, s.lo);
  s.lo += t.lo;
  return quick_two_sum(s.hi, s.lo);
}
__host__ __device__ __forceinline__ dd dd_neg(dd a) { return {-a.hi, -a.lo}; }
__host__ __device__ __forceinline__ dd dd_sub(dd a, dd b) { return dd_add(a, dd_neg(b)); }
__host__ __device__ __forceinline__ dd dd_mul(dd a, dd b) {
  dd p = two_prod(a.hi, b.hi);
  p.lo += a.hi * b.lo + a.lo * b.hi;
  return quick_two_sum(p.hi, p.lo);
}
__host__ __device__ __forceinline__ dd dd_mul_d(dd a, double b) {
  dd p = two_prod(a.hi, b);
  p.lo += a.lo * b;
  return quick_two_sum(p.hi, p.lo);
}
__host__ __device__ __forceinline__ dd dd_div(dd a, dd b) {
  double q1 = a.hi / b.hi;
  dd r = dd_sub(a, dd_mul_d(b, q1));
  double q2 = r.hi / b.hi;
  r = dd_sub(r, dd_mul_d(b, q2));
  double q3 = r.hi / b.hi;
  dd q = quick_two_sum(q1, q2);
  return dd_add(q, dd{q3, 0.0});
}
__host__ __device__ __forceinline__ dd dd_sqrt(dd a) {
  // one Newton step on top of the double square root (Karp's trick); a.hi > 0 is checked by the caller
  double x = 1.0 / sqrt(a.hi);
  double ax = a.hi * x;
  dd err = dd_sub(a, two_prod(ax, ax));
  return dd_add(dd{ax, 0.0}, dd{err.hi * (x * 0.5), 0.0});
}

// x^p for a small positive integer p, correctly rounded (the exact product is formed in double-double and rounded once).
__host__ __device__ __forceinline__ double pow_int_rounded(double x, int p) {
  if (p == 1) return x;
  dd r = two_prod(x, x);
  for (int i = 2; i < p; ++i) r = dd_mul_d(r, x);
  return r.hi + r.lo;
}

}  // namespace lsmc
