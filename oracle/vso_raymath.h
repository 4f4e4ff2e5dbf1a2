/*
 * vso_raymath.h -- TEST INFRASTRUCTURE ONLY (part of the CPU oracle, see vsnray_oracle.h).
 *
 * sin / cos / pow for the values that DEFINE A RAY in the sampling code of the reference
 * (cosine_sample_hemisphere, sampling.h:61-71; blinn::sample_f, brdf.h:121-150).  The
 * reference calls libm (std::sin / std::cos / std::pow on float); CUDA's sinf / cosf / powf
 * differ from libm by a few ulps, which forks a path at a triangle edge once in a few
 * thousand bounces.  Here the functions are evaluated in double with a FIXED sequence of
 * IEEE-754 additions, multiplications and divisions (each a single rounding; the file is
 * compiled with -ffp-contract=off) and rounded to float once.  csrc/shade.cuh restates the
 * same sequence with __dadd_rn / __dmul_rn / __ddiv_rn, so device and oracle produce the
 * same bits; against glibc they are the correctly rounded value except when the exact
 * result lies within ~1e-13 ulp(float) ... 2^-22 of a rounding boundary (glibc's own sinf /
 * cosf are correctly rounded there too in all but ~1e-8 of arguments; its powf has up to
 * 0.52 ulp of error, i.e. it differs from the correctly rounded value in about 1 % of
 * arguments, by one ulp).  tests/test_oracle.py counts the disagreements.
 */
#ifndef VSO_RAYMATH_H
#define VSO_RAYMATH_H

#include <math.h>
#include <stdint.h>
#include <string.h>

/* pi/2 split: HI has 33 significant bits, so k * HI is exact for |k| < 2^20 */
#define VSO_PIO2_HI 1.5707963267341256e+00
#define VSO_PIO2_LO 6.0771005065061922e-11
#define VSO_TWO_OVER_PI 6.3661977236758138e-01
/* ln 2 split: HI has 32 significant bits (k * HI exact for |k| < 2^21) */
#define VSO_LN2_HI 6.9314718036912382e-01
#define VSO_LN2_LO 1.9082149292705877e-10
#define VSO_INV_LN2 1.4426950408889634e+00

/* sin and cos of a float angle |x| < 2^20, correctly rounded to float (see header) */
static inline void vso_sincosf(float x, float* s, float* c)
{
    const double xd = (double)x;
    const double kd = rint(xd * VSO_TWO_OVER_PI);
    const double r = (xd - kd * VSO_PIO2_HI) - kd * VSO_PIO2_LO;     /* |r| <= pi/4 (+ 1 ulp) */
    const double r2 = r * r;
    /* Taylor series, Horner in r^2: sin to r^17, cos to r^18 (remainder < 5e-17 on |r| <= 0.786) */
    double ps = -1.0 / 355687428096000.0;                            /* -1/17! */
    ps = ps * r2 + 1.0 / 1307674368000.0;                            /* 1/15! */
    ps = ps * r2 - 1.0 / 6227020800.0;
    ps = ps * r2 + 1.0 / 39916800.0;
    ps = ps * r2 - 1.0 / 362880.0;
    ps = ps * r2 + 1.0 / 5040.0;
    ps = ps * r2 - 1.0 / 120.0;
    ps = ps * r2 + 1.0 / 6.0;
    const double sn = r - (r * r2) * ps;
    double pc = 1.0 / 6402373705728000.0;                            /* 1/18! */
    pc = pc * r2 - 1.0 / 20922789888000.0;                           /* -1/16! */
    pc = pc * r2 + 1.0 / 87178291200.0;
    pc = pc * r2 - 1.0 / 479001600.0;
    pc = pc * r2 + 1.0 / 3628800.0;
    pc = pc * r2 - 1.0 / 40320.0;
    pc = pc * r2 + 1.0 / 720.0;
    pc = pc * r2 - 1.0 / 24.0;
    pc = pc * r2 + 0.5;
    const double cs = 1.0 - r2 * pc;
    const int q = (int)((long long)kd & 3);
    const double so = (q & 1) ? cs : sn, co = (q & 1) ? sn : cs;
    *s = (float)((q & 2) ? -so : so);
    *c = (float)(((q + 1) & 2) ? -co : co);
}

/* x^y for x >= 0, finite y, rounded to float once; x < 0 -> NaN (never asked for on this path) */
static inline float vso_powf(float x, float y)
{
    if (y == 0.0f) return 1.0f;
    if (x != x || y != y || x < 0.0f) return NAN;
    if (x == 0.0f) return y > 0.0f ? 0.0f : INFINITY;
    if (x == 1.0f) return 1.0f;
    if (x == INFINITY) return y > 0.0f ? INFINITY : 0.0f;
    const double xd = (double)x;                                     /* float subnormals are normal doubles */
    uint64_t b; memcpy(&b, &xd, 8);
    int k = (int)((b >> 52) & 0x7ff) - 1023;
    b = (b & 0x000fffffffffffffull) | 0x3ff0000000000000ull;
    double m; memcpy(&m, &b, 8);                                     /* [1, 2) */
    if (m > 1.4142135623730951) { m = m * 0.5; k = k + 1; }         /* [sqrt(1/2), sqrt(2)] */
    const double s = (m - 1.0) / (m + 1.0);
    const double s2 = s * s;                                         /* <= 0.02944 */
    /* ln m = 2 atanh(s) = 2 s (1 + s2/3 + s2^2/5 + ... + s2^10/21); remainder < 1e-18 */
    double p = 1.0 / 21.0;
    p = p * s2 + 1.0 / 19.0;
    p = p * s2 + 1.0 / 17.0;
    p = p * s2 + 1.0 / 15.0;
    p = p * s2 + 1.0 / 13.0;
    p = p * s2 + 1.0 / 11.0;
    p = p * s2 + 1.0 / 9.0;
    p = p * s2 + 1.0 / 7.0;
    p = p * s2 + 1.0 / 5.0;
    p = p * s2 + 1.0 / 3.0;
    const double lnm = 2.0 * s + (2.0 * s) * (s2 * p);
    const double kd = (double)k;
    const double lnx = kd * VSO_LN2_HI + (kd * VSO_LN2_LO + lnm);
    const double z = (double)y * lnx;
    if (z > 89.0) return INFINITY;
    if (z < -104.0) return 0.0f;
    const double nd = rint(z * VSO_INV_LN2);
    const double r = (z - nd * VSO_LN2_HI) - nd * VSO_LN2_LO;        /* |r| <= 0.3466 */
    double e = 1.0 / 6227020800.0;                                   /* 1/13! */
    e = e * r + 1.0 / 479001600.0;
    e = e * r + 1.0 / 39916800.0;
    e = e * r + 1.0 / 3628800.0;
    e = e * r + 1.0 / 362880.0;
    e = e * r + 1.0 / 40320.0;
    e = e * r + 1.0 / 5040.0;
    e = e * r + 1.0 / 720.0;
    e = e * r + 1.0 / 120.0;
    e = e * r + 1.0 / 24.0;
    e = e * r + 1.0 / 6.0;
    e = e * r + 0.5;
    e = e * r + 1.0;
    e = e * r + 1.0;
    const uint64_t sb = (uint64_t)((long long)nd + 1023) << 52;      /* 2^n, n in [-151, 129] */
    double sc; memcpy(&sc, &sb, 8);
    return (float)(e * sc);
}

#endif
