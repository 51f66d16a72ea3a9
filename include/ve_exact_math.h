/* ve_exact_math.h — transcendental functions that give BIT-IDENTICAL results under
 * g++ (host, -ffp-contract=off, hardware FMA) and nvcc (device, --fmad=false).
 *
 * Why: the erosion path calls f64 `powf/exp/ln/ln_1p/tanh` and f32 `powf`
 * (reference call sites: world/src/sim/erosion.rs:571-575, 903-904, 926, 940, 988-989,
 * 996-997, 1343-1347, 1524, 2468).  Rust lowers those to the host libm; CUDA's libdevice
 * rounds differently in the last place.  Several results are cast to f32 or compared
 * against thresholds, so a 1-ulp difference can flip discrete decisions.  Both the CUDA
 * kernels and the CPU oracle therefore evaluate the SAME sequence of IEEE-754 operations
 * (+ - * / sqrt and explicitly written fma), which is what this header provides.
 *
 * Accuracy: double-double internals, error < 0.51 ulp for log/exp/pow/log1p and < 1 ulp
 * for tanh on the ranges the path uses; measured against glibc in
 * tests/test_exact_math.py (the measured deviation is the "oracle vs host libm" figure
 * DESIGN.md quotes).  Not a general libm: arguments are assumed finite; pow requires
 * x >= 0.
 *
 * Algorithms are the textbook ones (atanh series for log, Taylor series for exp after
 * ln2 range reduction) written from scratch for this project.
 */
#ifndef VE_EXACT_MATH_H
#define VE_EXACT_MATH_H

#include <stdint.h>
#include <string.h>
#include <math.h>

#if defined(__CUDACC__)
#define VM_HD __host__ __device__ __forceinline__
#else
#define VM_HD static inline
#endif

typedef struct { double hi, lo; } vm_dd;

VM_HD double vm_fma(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return __builtin_fma(a, b, c);
#endif
}
VM_HD double vm_sqrt(double x) {
#if defined(__CUDA_ARCH__)
    return __dsqrt_rn(x);
#else
    return __builtin_sqrt(x);
#endif
}
VM_HD float vm_sqrtf(float x) {
#if defined(__CUDA_ARCH__)
    return __fsqrt_rn(x);
#else
    return __builtin_sqrtf(x);
#endif
}
VM_HD uint64_t vm_bits(double x) {
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(x);
#else
    uint64_t u; memcpy(&u, &x, 8); return u;
#endif
}
VM_HD double vm_from_bits(uint64_t u) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double x; memcpy(&x, &u, 8); return x;
#endif
}

/* error-free transforms */
VM_HD vm_dd vm_two_sum(double a, double b) {
    vm_dd r; r.hi = a + b; double bb = r.hi - a; r.lo = (a - (r.hi - bb)) + (b - bb); return r;
}
VM_HD vm_dd vm_fast_two_sum(double a, double b) { /* |a| >= |b| */
    vm_dd r; r.hi = a + b; r.lo = b - (r.hi - a); return r;
}
VM_HD vm_dd vm_two_prod(double a, double b) {
    vm_dd r; r.hi = a * b; r.lo = vm_fma(a, b, -r.hi); return r;
}
VM_HD vm_dd vm_dd_add(vm_dd a, vm_dd b) {
    vm_dd s = vm_two_sum(a.hi, b.hi);
    double t = (a.lo + b.lo) + s.lo;
    return vm_fast_two_sum(s.hi, t);
}
VM_HD vm_dd vm_dd_add_d(vm_dd a, double b) {
    vm_dd s = vm_two_sum(a.hi, b);
    double t = a.lo + s.lo;
    return vm_fast_two_sum(s.hi, t);
}
VM_HD vm_dd vm_dd_mul(vm_dd a, vm_dd b) {
    vm_dd p = vm_two_prod(a.hi, b.hi);
    double t = (a.hi * b.lo + a.lo * b.hi) + p.lo;
    return vm_fast_two_sum(p.hi, t);
}
VM_HD vm_dd vm_dd_mul_d(vm_dd a, double b) {
    vm_dd p = vm_two_prod(a.hi, b);
    double t = a.lo * b + p.lo;
    return vm_fast_two_sum(p.hi, t);
}

/* log(x) for finite x > 0 as a double-double. */
VM_HD vm_dd vm_log_dd(double x) {
    int e = 0;
    uint64_t u = vm_bits(x);
    if ((u >> 52) == 0) { /* subnormal: scale by 2^54 */
        x = x * 18014398509481984.0; u = vm_bits(x); e = -54;
    }
    e += (int)((u >> 52) & 0x7ff) - 1023;
    u = (u & 0x000fffffffffffffULL) | 0x3ff0000000000000ULL;
    double m = vm_from_bits(u); /* [1,2) */
    if (m > 1.4142135623730951) { m = m * 0.5; e += 1; }
    /* s = (m-1)/(m+1) in double-double */
    double num = m - 1.0;                 /* exact */
    vm_dd den = vm_two_sum(m, 1.0);       /* exact as a pair */
    double s_hi = num / den.hi;
    double rem = vm_fma(-s_hi, den.hi, num);
    double s_lo = (rem - s_hi * den.lo) / den.hi;
    vm_dd s; s.hi = s_hi; s.lo = s_lo;
    /* z = s^2, s3 = s^3 in double-double */
    vm_dd z = vm_dd_mul(s, s);
    vm_dd s3 = vm_dd_mul(z, s);
    /* tail Q(z) = 1/5 + z/7 + ... + z^11/27, plain double */
    double zh = z.hi;
    double q = 1.0 / 27.0;
    q = q * zh + 1.0 / 25.0;
    q = q * zh + 1.0 / 23.0;
    q = q * zh + 1.0 / 21.0;
    q = q * zh + 1.0 / 19.0;
    q = q * zh + 1.0 / 17.0;
    q = q * zh + 1.0 / 15.0;
    q = q * zh + 1.0 / 13.0;
    q = q * zh + 1.0 / 11.0;
    q = q * zh + 1.0 / 9.0;
    q = q * zh + 1.0 / 7.0;
    q = q * zh + 1.0 / 5.0;
    double s5q = (s3.hi * zh) * q;        /* s^5 * Q */
    /* third = 1/3 as double-double */
    vm_dd third; third.hi = 0.33333333333333331; third.lo = 1.8503717077085941e-17;
    vm_dd t3 = vm_dd_mul(s3, third);
    vm_dd acc = vm_dd_add_d(t3, s5q);
    acc = vm_dd_add(s, acc);              /* s + s^3/3 + s^5 Q */
    acc.hi = acc.hi * 2.0; acc.lo = acc.lo * 2.0; /* log(m) */
    /* + e*ln2 */
    vm_dd ln2; ln2.hi = 0.69314718055994529; ln2.lo = 2.3190468138462996e-17;
    vm_dd el = vm_dd_mul_d(ln2, (double)e);
    return vm_dd_add(el, acc);
}

VM_HD double vm_log(double x) {
    if (x == 1.0) return 0.0;
    vm_dd r = vm_log_dd(x);
    return r.hi + r.lo;
}

/* log(1+x) for x > -1 (reference use: `(-x).ln_1p()` with x in (0,1), erosion.rs:571). */
VM_HD double vm_log1p(double x) {
    vm_dd u = vm_two_sum(1.0, x);
    if (u.hi == 1.0 && u.lo == 0.0) return x;
    vm_dd r = vm_log_dd(u.hi);
    double corr = u.lo / u.hi;
    r = vm_dd_add_d(r, corr);
    return r.hi + r.lo;
}

/* exp(p.hi + p.lo) - returns T and k such that exp(p) = 2^k * (1 + T), T double-double. */
VM_HD vm_dd vm_expm1_reduced(vm_dd p, int* kout) {
    double kf = p.hi * 1.4426950408889634;
    /* round to nearest integer without relying on rint() */
    double kr = (kf >= 0.0) ? (double)(long long)(kf + 0.5) : (double)(long long)(kf - 0.5);
    int k = (int)kr;
    vm_dd ln2; ln2.hi = 0.69314718055994529; ln2.lo = 2.3190468138462996e-17;
    vm_dd kl = vm_dd_mul_d(ln2, -kr);
    vm_dd r = vm_dd_add(p, kl);           /* |r| <= ~0.35 */
    double rh = r.hi;
    /* R4(r) = 1/24 + r/120 + ... + r^11/15! */
    double c = 1.0 / 1307674368000.0;
    c = c * rh + 1.0 / 87178291200.0;
    c = c * rh + 1.0 / 6227020800.0;
    c = c * rh + 1.0 / 479001600.0;
    c = c * rh + 1.0 / 39916800.0;
    c = c * rh + 1.0 / 3628800.0;
    c = c * rh + 1.0 / 362880.0;
    c = c * rh + 1.0 / 40320.0;
    c = c * rh + 1.0 / 5040.0;
    c = c * rh + 1.0 / 720.0;
    c = c * rh + 1.0 / 120.0;
    c = c * rh + 1.0 / 24.0;
    double rr = rh * rh;
    double r4c = (rr * rr) * c;
    vm_dd r2 = vm_dd_mul(r, r);
    r2.hi = r2.hi * 0.5; r2.lo = r2.lo * 0.5;            /* r^2/2 */
    vm_dd third; third.hi = 0.33333333333333331; third.lo = 1.8503717077085941e-17;
    vm_dd r3 = vm_dd_mul(vm_dd_mul(r2, r), third);       /* r^3/6 */
    vm_dd t = vm_dd_add_d(r3, r4c);
    t = vm_dd_add(r2, t);
    t = vm_dd_add(r, t);                  /* T = r + r^2/2 + r^3/6 + r^4 R4 */
    *kout = k;
    return t;
}

VM_HD double vm_ldexp_norm(double x, int k) {
    /* x * 2^k, exact while the result stays normal; two-step for the extremes */
    if (k > 1000) { x = x * vm_from_bits(0x7fe0000000000000ULL); k -= 1023; }   /* 2^1023 */
    if (k < -1000) { x = x * vm_from_bits(0x0360000000000000ULL); k += 969; }   /* 2^-969 */
    return x * vm_from_bits((uint64_t)(k + 1023) << 52);
}

VM_HD double vm_exp_dd(vm_dd p) {
    if (p.hi > 709.78) return INFINITY;
    if (p.hi < -745.2) return 0.0;
    int k;
    vm_dd t = vm_expm1_reduced(p, &k);
    vm_dd one_t = vm_dd_add_d(t, 1.0);
    return vm_ldexp_norm(one_t.hi + one_t.lo, k);
}

VM_HD double vm_exp(double x) {
    vm_dd p; p.hi = x; p.lo = 0.0;
    return vm_exp_dd(p);
}

/* pow(x, y) for x >= 0 (Rust f64::powf call sites listed at the top). */
VM_HD double vm_pow(double x, double y) {
    if (y == 0.0) return 1.0;
    if (x == 1.0) return 1.0;
    if (y == 1.0) return x;
    if (x == 0.0) return (y > 0.0) ? 0.0 : INFINITY;
    vm_dd l = vm_log_dd(x);
    vm_dd p = vm_dd_mul_d(l, y);
    return vm_exp_dd(p);
}

/* f32 powf (reference: SimdType = f32 Newton loop, erosion.rs:1343-1347).  y == 0 -> 1
 * (any x, including 0), y == 0.5 -> correctly rounded sqrtf (what LLVM lowers a constant
 * 0.5 exponent to); otherwise evaluated in f64 and rounded once. */
VM_HD float vm_powf(float x, float y) {
    if (y == 0.0f) return 1.0f;
    if (y == 1.0f) return x;
    if (y == 0.5f) return vm_sqrtf(x);
    return (float)vm_pow((double)x, (double)y);
}

/* tanh(x) (reference: logistic_cdf, erosion.rs:575). */
VM_HD double vm_tanh(double x) {
    double ax = x < 0.0 ? -x : x;
    if (ax < 3.7252902984619141e-09) return x;          /* 2^-28 */
    if (ax > 22.0) return x < 0.0 ? -1.0 : 1.0;
    vm_dd p; p.hi = 2.0 * ax; p.lo = 0.0;
    int k;
    vm_dd t = vm_expm1_reduced(p, &k);                   /* e^{2|x|} = 2^k (1+T) */
    double sc = vm_ldexp_norm(1.0, k);
    vm_dd em1; em1.hi = t.hi * sc; em1.lo = t.lo * sc;   /* 2^k T (exact scaling) */
    em1 = vm_dd_add_d(em1, sc - 1.0);                    /* + (2^k - 1), exact for k<=52 */
    vm_dd den = vm_dd_add_d(em1, 2.0);
    /* q = em1/den in double-double, then round */
    double q_hi = em1.hi / den.hi;
    double rem = vm_fma(-q_hi, den.hi, em1.hi);
    double q_lo = ((rem + em1.lo) - q_hi * den.lo) / den.hi;
    double r = q_hi + q_lo;
    return x < 0.0 ? -r : r;
}

#endif /* VE_EXACT_MATH_H */
