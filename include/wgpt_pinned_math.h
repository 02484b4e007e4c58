/*
 * wgpt_pinned_math.h -- pinned transcendental functions for the path-tracing hot path.
 *
 * WGSL leaves the precision of sin/cos/acos/log/pow loose, and glibc and CUDA libdevice
 * disagree by ulps.  To make WHOLE PATHS bit-comparable between the CPU oracle and the
 * exact-mode CUDA kernel, both sides compile THIS header: every function is evaluated in
 * IEEE binary64 with plain + - * / sqrt (never contracted into FMA: build with
 * `-ffp-contract=off` / `-fmad=false`) and rounded once to binary32.  The binary64 result is
 * accurate to ~1e-15, so the float result is the correctly rounded one except in ~1e-8 of
 * calls; tests/test_pinned_math.py checks them against glibc's float functions.
 *
 * Call sites in the reference (file:line under /root/reference):
 *   sin/cos : shaders/importanceSampling.wgsl:11-13,41-42  shaders/scatterRay.wgsl:84
 *   acos    : shaders/importanceSampling.wgsl:9
 *   log     : shaders/common.wgsl:134
 *   pow     : shaders/importanceSampling.wgsl:4  shaders/scatterRay.wgsl:80  shaders/fragment.js:28
 *
 * Header-only, C++ (host and CUDA device).  No state, no tables.
 */
#ifndef WGPT_PINNED_MATH_H
#define WGPT_PINNED_MATH_H

#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define WGPT_HD __host__ __device__ __forceinline__
#else
#define WGPT_HD static inline
#endif

namespace wgpt_pm {

/* multiply / add / subtract that can never be fused into an FMA: on the device the _rn
 * intrinsics (ptxas never contracts them, whatever -fmad says); on the host plain operators,
 * which the oracle's build pins with -ffp-contract=off. */
#if defined(__CUDA_ARCH__)
WGPT_HD double dm(double a, double b) { return __dmul_rn(a, b); }
WGPT_HD double da(double a, double b) { return __dadd_rn(a, b); }
WGPT_HD double ds(double a, double b) { return __dsub_rn(a, b); }
#else
WGPT_HD double dm(double a, double b) { return a * b; }
WGPT_HD double da(double a, double b) { return a + b; }
WGPT_HD double ds(double a, double b) { return a - b; }
#endif
/* Horner step p*z + c */
WGPT_HD double hs(double p, double z, double c) { return da(dm(p, z), c); }

WGPT_HD int64_t d2bits(double d) { int64_t b; memcpy(&b, &d, 8); return b; }
WGPT_HD double  bits2d(int64_t b) { double d; memcpy(&d, &b, 8); return d; }

/* sin and cos of r, |r| <= pi/4 (Taylor, Horner in r*r; remainder < 5e-17). */
WGPT_HD double ksin(double r) {
    const double z = dm(r, r);
    double p = 2.8114572543455206e-15;            /*  1/17! */
    p = hs(p, z, -7.6471637318198164e-13);          /* -1/15! */
    p = hs(p, z, 1.6059043836821613e-10);           /*  1/13! */
    p = hs(p, z, -2.5052108385441720e-08);          /* -1/11! */
    p = hs(p, z, 2.7557319223985893e-06);           /*  1/9!  */
    p = hs(p, z, -1.9841269841269841e-04);          /* -1/7!  */
    p = hs(p, z, 8.3333333333333332e-03);           /*  1/5!  */
    p = hs(p, z, -1.6666666666666666e-01);          /* -1/3!  */
    return da(r, dm(r, dm(z, p)));
}
WGPT_HD double kcos(double r) {
    const double z = dm(r, r);
    double p = 4.7794773323873853e-14;            /*  1/16! */
    p = hs(p, z, -1.1470745597729725e-11);          /* -1/14! */
    p = hs(p, z, 2.0876756987868100e-09);           /*  1/12! */
    p = hs(p, z, -2.7557319223985888e-07);          /* -1/10! */
    p = hs(p, z, 2.4801587301587302e-05);           /*  1/8!  */
    p = hs(p, z, -1.3888888888888889e-03);          /* -1/6!  */
    p = hs(p, z, 4.1666666666666664e-02);           /*  1/4!  */
    p = hs(p, z, -0.5);                             /* -1/2!  */
    return da(1.0, dm(z, p));
}

/* Two-part Cody-Waite reduction by pi/2 (k*PIO2_HI is exact for |k| < 2^20). */
WGPT_HD double reduce_pio2(double x, int* quadrant) {
    const double PIO2_HI = 1.57079632673412561417e+00;   /* 0x3FF921FB54400000 */
    const double PIO2_LO = 6.07710050650619224932e-11;   /* 0x3DD0B4611A626331 */
    const double k = rint(dm(x, 6.36619772367581382433e-01));
    *quadrant = (int)((long long)k & 3);
    return ds(ds(x, dm(k, PIO2_HI)), dm(k, PIO2_LO));
}

WGPT_HD double sin_d(double x) {
    int q; const double r = reduce_pio2(x, &q);
    switch (q) { case 0: return ksin(r); case 1: return kcos(r); case 2: return -ksin(r); default: return -kcos(r); }
}
WGPT_HD double cos_d(double x) {
    int q; const double r = reduce_pio2(x, &q);
    switch (q) { case 0: return kcos(r); case 1: return -ksin(r); case 2: return -kcos(r); default: return ksin(r); }
}

/* asin for |z| <= 0.5: series start + two Newton steps on sin (cos >= 0.866, well conditioned). */
WGPT_HD double kasin(double z) {
    const double z2 = dm(z, z);
    double p = 15.0 / 336.0;
    p = hs(p, z2, 3.0 / 40.0);
    p = hs(p, z2, 1.0 / 6.0);
    p = hs(p, z2, 1.0);
    double t = dm(z, p);
    t = ds(t, ds(ksin(t), z) / kcos(t));
    t = ds(t, ds(ksin(t), z) / kcos(t));
    t = ds(t, ds(ksin(t), z) / kcos(t));
    return t;
}
WGPT_HD double acos_d(double x) {
    const double PI_D = 3.14159265358979311600e+00, PIO2_D = 1.57079632679489655800e+00;
    if (x != x || x > 1.0 || x < -1.0) return bits2d(0x7ff8000000000000LL);
    if (x > 0.5)  return dm(2.0, kasin(sqrt(dm(ds(1.0, x), 0.5))));
    if (x < -0.5) return ds(PI_D, dm(2.0, kasin(sqrt(dm(da(1.0, x), 0.5)))));
    return ds(PIO2_D, kasin(x));
}

/* natural log of a positive finite normal double (floats widen to normal doubles). */
WGPT_HD double log_pos(double d) {
    int64_t b = d2bits(d);
    int e = (int)((b >> 52) & 0x7ff) - 1023;
    double m = bits2d((b & 0x000fffffffffffffLL) | 0x3ff0000000000000LL);
    if (m > 1.41421356237309514547) { m = dm(m, 0.5); e += 1; }
    const double s = ds(m, 1.0) / da(m, 1.0), w = dm(s, s);
    double p = 1.0 / 23.0;
    p = hs(p, w, 1.0 / 21.0);  p = hs(p, w, 1.0 / 19.0);  p = hs(p, w, 1.0 / 17.0);
    p = hs(p, w, 1.0 / 15.0);  p = hs(p, w, 1.0 / 13.0);  p = hs(p, w, 1.0 / 11.0);
    p = hs(p, w, 1.0 / 9.0);   p = hs(p, w, 1.0 / 7.0);   p = hs(p, w, 1.0 / 5.0);
    p = hs(p, w, 1.0 / 3.0);
    const double s2 = dm(2.0, s);
    const double logm = da(s2, dm(s2, dm(w, p)));
    return da(dm((double)e, 6.93147180559945286227e-01), logm);
}
WGPT_HD double log_d(double x) {
    if (x != x || x < 0.0) return bits2d(0x7ff8000000000000LL);
    if (x == 0.0) return bits2d((int64_t)0xfff0000000000000ULL);   /* -inf */
    if (x > 1.7e308) return x;                                   /* +inf */
    return log_pos(x);
}

WGPT_HD double exp_d(double t) {
    if (t != t) return t;
    if (t > 700.0) return bits2d(0x7ff0000000000000LL);
    if (t < -700.0) return 0.0;
    const double LN2_HI = 6.93147180369123816490e-01;    /* 0x3fe62e42fee00000 */
    const double LN2_LO = 1.90821492927058770002e-10;    /* 0x3dea39ef35793c76 */
    const double k = rint(dm(t, 1.44269504088896338700e+00));
    const double r = ds(ds(t, dm(k, LN2_HI)), dm(k, LN2_LO));
    double p = 1.6059043836821613e-10;            /* 1/13! */
    p = hs(p, r, 2.0876756987868100e-09);           /* 1/12! */
    p = hs(p, r, 2.5052108385441720e-08);           /* 1/11! */
    p = hs(p, r, 2.7557319223985888e-07);           /* 1/10! */
    p = hs(p, r, 2.7557319223985893e-06);           /* 1/9!  */
    p = hs(p, r, 2.4801587301587302e-05);           /* 1/8!  */
    p = hs(p, r, 1.9841269841269841e-04);           /* 1/7!  */
    p = hs(p, r, 1.3888888888888889e-03);           /* 1/6!  */
    p = hs(p, r, 8.3333333333333332e-03);           /* 1/5!  */
    p = hs(p, r, 4.1666666666666664e-02);           /* 1/4!  */
    p = hs(p, r, 1.6666666666666666e-01);           /* 1/3!  */
    p = hs(p, r, 0.5);
    p = hs(p, r, 1.0);
    p = hs(p, r, 1.0);
    return dm(p, bits2d(((int64_t)k + 1023) << 52));
}

/* ---- the float entry points the shader restatements call ---- */
WGPT_HD float sinf_(float x)  { return (float)sin_d((double)x); }
WGPT_HD float cosf_(float x)  { return (float)cos_d((double)x); }
/* sinf_(x) and cosf_(x) at once: one reduction, both kernels evaluated by every caller (no
 * quadrant-dependent control flow, which would split a GPU warp four ways); every lane performs
 * the same operations as sin_d / cos_d on the same inputs, so the bits equal the separate calls. */
WGPT_HD void sincosf_(float x, float* s, float* c) {
    int q; const double r = reduce_pio2((double)x, &q);
    const double ks = ksin(r), kc = kcos(r);
    const double sv = (q & 1) ? kc : ks, cv = (q & 1) ? ks : kc;
    *s = (float)((q & 2) ? -sv : sv);
    *c = (float)(((q + 1) & 2) ? -cv : cv);
}
WGPT_HD float acosf_(float x) { return (float)acos_d((double)x); }
WGPT_HD float logf_(float x)  { return (float)log_d((double)x); }
/* WGSL pow(x,y): defined for x >= 0 here; x < 0 -> NaN (the WGSL spec leaves it undefined). */
WGPT_HD float powf_(float x, float y) {
    if (x != x || y != y) return x + y;
    if (y == 0.0f || x == 1.0f) return 1.0f;
    if (x < 0.0f) return (float)bits2d(0x7ff8000000000000LL);
    if (x == 0.0f) return y > 0.0f ? 0.0f : (float)bits2d(0x7ff0000000000000LL);
    return (float)exp_d(dm((double)y, log_d((double)x)));
}

}  /* namespace wgpt_pm */
#endif
