/*
 * ag2_detmath.h -- exactly reproducible log / cos for the importance sampler's Gaussian draws.
 *
 * The reference draws its samples with boost::normal_distribution on a time-seeded boost::mt19937
 * (importance_sampling.cpp:57-60,120,134): not reproducible from run to run.  Deviation D5 replaces
 * it by Box-Muller on a counter-based stream.  For the candidate generator to run on the GPU and
 * still give the oracle's candidates bit for bit, the two transcendental functions cannot be
 * libm's or CUDA's (they differ in the last ulp): they are defined HERE, as fixed sequences of IEEE
 * double multiplications, additions, one division and bit operations -- every operation correctly
 * rounded on both sides, no fused multiply-add (explicit __dmul_rn / __dadd_rn on the device; the
 * oracle is built -ffp-contract=off).  The same source is compiled into libag2.so (importance.cu)
 * and into the CPU oracle (oracle/ag2_oracle.cpp), which is what makes the draws identical.
 * Accuracy is a few ulp, which is all a sampler needs.
 */
#ifndef AG2_DETMATH_H_
#define AG2_DETMATH_H_

#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define AG2_DM_FN __host__ __device__ __forceinline__
#else
#define AG2_DM_FN static inline
#endif

AG2_DM_FN double ag2_dm_mul(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dmul_rn(a, b);
#else
  volatile double r = a * b; /* volatile: the product is rounded to double before any use */
  return r;
#endif
}
AG2_DM_FN double ag2_dm_add(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dadd_rn(a, b);
#else
  volatile double r = a + b;
  return r;
#endif
}
AG2_DM_FN double ag2_dm_div(double a, double b) {
#ifdef __CUDA_ARCH__
  return __ddiv_rn(a, b);
#else
  volatile double r = a / b;
  return r;
#endif
}
AG2_DM_FN uint64_t ag2_dm_bits(double x) {
#ifdef __CUDA_ARCH__
  return (uint64_t)__double_as_longlong(x);
#else
  uint64_t u;
  memcpy(&u, &x, 8);
  return u;
#endif
}
AG2_DM_FN double ag2_dm_from_bits(uint64_t u) {
#ifdef __CUDA_ARCH__
  return __longlong_as_double((long long)u);
#else
  double x;
  memcpy(&x, &u, 8);
  return x;
#endif
}

/* Natural logarithm of a positive, finite, normal x.  x = m * 2^e with m in [sqrt(1/2), sqrt(2));
 * log m = 2 atanh(s), s = (m - 1) / (m + 1), |s| < 0.1716: odd series to s^23 by Horner in s^2. */
AG2_DM_FN double ag2_det_log(double x) {
  uint64_t u = ag2_dm_bits(x);
  int e = (int)((u >> 52) & 0x7FFu) - 1023;
  uint64_t frac = u & 0x000FFFFFFFFFFFFFull;
  double m = ag2_dm_from_bits(frac | 0x3FF0000000000000ull); /* [1, 2) */
  if (m > 1.4142135623730951) {                                /* -> [sqrt(1/2), sqrt(2)) */
    m = ag2_dm_mul(m, 0.5);                                    /* exact */
    e += 1;
  }
  const double s = ag2_dm_div(ag2_dm_add(m, -1.0), ag2_dm_add(m, 1.0));
  const double z = ag2_dm_mul(s, s);
  double p = 1.0 / 23.0;
  p = ag2_dm_add(ag2_dm_mul(p, z), 1.0 / 21.0);
  p = ag2_dm_add(ag2_dm_mul(p, z), 1.0 / 19.0);
  p = ag2_dm_add(ag2_dm_mul(p, z), 1.0 / 17.0);
  p = ag2_dm_add(ag2_dm_mul(p, z), 1.0 / 15.0);
  p = ag2_dm_add(ag2_dm_mul(p, z), 1.0 / 13.0);
  p = ag2_dm_add(ag2_dm_mul(p, z), 1.0 / 11.0);
  p = ag2_dm_add(ag2_dm_mul(p, z), 1.0 / 9.0);
  p = ag2_dm_add(ag2_dm_mul(p, z), 1.0 / 7.0);
  p = ag2_dm_add(ag2_dm_mul(p, z), 1.0 / 5.0);
  p = ag2_dm_add(ag2_dm_mul(p, z), 1.0 / 3.0);
  p = ag2_dm_add(ag2_dm_mul(p, z), 1.0);
  const double logm = ag2_dm_mul(ag2_dm_mul(2.0, s), p);
  return ag2_dm_add(ag2_dm_mul((double)e, 0.6931471805599453), logm);
}

/* cos(2 pi u) for u in [0, 1].  k = nearest quarter turn, r = u - k/4 (exact), |r| <= 1/8;
 * cos(2 pi u) = cos(k pi/2 + z), z = 2 pi r, |z| <= pi/4: Taylor polynomials of cos / sin in z^2. */
AG2_DM_FN double ag2_det_cos2pi(double u) {
  const int k = (int)ag2_dm_add(ag2_dm_mul(u, 4.0), 0.5); /* 0 .. 4 */
  const double r = ag2_dm_add(u, -0.25 * (double)k);      /* exact: both are multiples of 2^-55 below 2 */
  const double z = ag2_dm_mul(6.283185307179586, r);
  const double w = ag2_dm_mul(z, z);
  double c = 1.0 / 20922789888000.0; /* 1/16! */
  c = ag2_dm_add(ag2_dm_mul(c, w), -1.0 / 87178291200.0);
  c = ag2_dm_add(ag2_dm_mul(c, w), 1.0 / 479001600.0);
  c = ag2_dm_add(ag2_dm_mul(c, w), -1.0 / 3628800.0);
  c = ag2_dm_add(ag2_dm_mul(c, w), 1.0 / 40320.0);
  c = ag2_dm_add(ag2_dm_mul(c, w), -1.0 / 720.0);
  c = ag2_dm_add(ag2_dm_mul(c, w), 1.0 / 24.0);
  c = ag2_dm_add(ag2_dm_mul(c, w), -0.5);
  c = ag2_dm_add(ag2_dm_mul(c, w), 1.0);
  double sn = -1.0 / 355687428096000.0; /* -1/17! */
  sn = ag2_dm_add(ag2_dm_mul(sn, w), 1.0 / 1307674368000.0);
  sn = ag2_dm_add(ag2_dm_mul(sn, w), -1.0 / 6227020800.0);
  sn = ag2_dm_add(ag2_dm_mul(sn, w), 1.0 / 39916800.0);
  sn = ag2_dm_add(ag2_dm_mul(sn, w), -1.0 / 362880.0);
  sn = ag2_dm_add(ag2_dm_mul(sn, w), 1.0 / 5040.0);
  sn = ag2_dm_add(ag2_dm_mul(sn, w), -1.0 / 120.0);
  sn = ag2_dm_add(ag2_dm_mul(sn, w), 1.0 / 6.0);
  sn = ag2_dm_add(ag2_dm_mul(sn, w), -1.0); /* sn = -(sin z) / z ... */
  sn = ag2_dm_mul(sn, z);                   /* ... = -sin z */
  switch (k & 3) {
    case 0: return c;
    case 1: return sn;   /* cos(pi/2 + z) = -sin z */
    case 2: return -c;
    default: return -sn; /* cos(3pi/2 + z) = sin z */
  }
}

/* One N(0,1) value from two uniforms in (0, 1): Box-Muller, sqrt(-2 ln u1) * cos(2 pi u2). */
AG2_DM_FN double ag2_det_normal(double u1, double u2) {
  const double t = ag2_dm_mul(-2.0, ag2_det_log(u1));
#ifdef __CUDA_ARCH__
  const double r = __dsqrt_rn(t);
#else
  volatile double r = __builtin_sqrt(t); /* IEEE sqrt: correctly rounded on both sides */
#endif
  return ag2_dm_mul(r, ag2_det_cos2pi(u2));
}

#endif /* AG2_DETMATH_H_ */
