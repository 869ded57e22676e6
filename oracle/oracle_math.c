/*
 * oracle_math.c -- TEST INFRASTRUCTURE ONLY.
 * fdlibm-style log / log10 / exp in strict IEEE double arithmetic.
 * Build: gcc -O2 -ffp-contract=off (see oracle/Makefile).
 *
 * The algorithms, evaluation order and coefficient tables follow fdlibm's e_log.c, e_log10.c and e_exp.c
 * (third-party, not part of the reference), whose notice is preserved here as it requires:
 *
 * ====================================================
 * Copyright (C) 1993 by Sun Microsystems, Inc. All rights reserved.
 *
 * Developed at SunSoft, a Sun Microsystems, Inc. business.
 * Permission to use, copy, modify, and distribute this
 * software is freely granted, provided that this notice
 * is preserved.
 * ====================================================
 */
#include "oracle_math.h"
#include <stdint.h>
#include <string.h>

static inline uint64_t d2u(double d) { uint64_t u; memcpy(&u, &d, 8); return u; }
static inline double   u2d(uint64_t u) { double d; memcpy(&d, &u, 8); return d; }
static inline int32_t  hi_word(double d) { return (int32_t)(d2u(d) >> 32); }
static inline uint32_t lo_word(double d) { return (uint32_t)d2u(d); }
static inline double   with_hi(double d, int32_t hi) {
    return u2d(((uint64_t)(uint32_t)hi << 32) | (d2u(d) & 0xffffffffu));
}

/* constants given by their bit patterns so no decimal parsing is involved */
#define C_LN2_HI   u2d(0x3fe62e42fee00000ull)
#define C_LN2_LO   u2d(0x3dea39ef35793c76ull)
#define C_TWO54    u2d(0x4350000000000000ull)
#define C_LG1      u2d(0x3FE5555555555593ull)
#define C_LG2      u2d(0x3FD999999997FA04ull)
#define C_LG3      u2d(0x3FD2492494229359ull)
#define C_LG4      u2d(0x3FCC71C51D8E78AFull)
#define C_LG5      u2d(0x3FC7466496CB03DEull)
#define C_LG6      u2d(0x3FC39A09D078C69Full)
#define C_LG7      u2d(0x3FC2F112DF3E5244ull)

double orc_log(double x)
{
    const double zero = 0.0;
    int32_t hx = hi_word(x);
    uint32_t lx = lo_word(x);
    int32_t k = 0, i, j;
    double f, s, z, w, t1, t2, R, hfsq, dk;

    if (hx < 0x00100000) {                       /* x < 2^-1022 */
        if (((hx & 0x7fffffff) | lx) == 0) return -C_TWO54 / zero;   /* log(0) = -inf */
        if (hx < 0) return (x - x) / zero;                          /* log(<0) = NaN  */
        k -= 54; x *= C_TWO54; hx = hi_word(x);
    }
    if (hx >= 0x7ff00000) return x + x;
    k += (hx >> 20) - 1023;
    hx &= 0x000fffff;
    i = (hx + 0x95f64) & 0x100000;
    x = with_hi(x, hx | (i ^ 0x3ff00000));       /* normalise x or x/2 */
    k += (i >> 20);
    f = x - 1.0;
    if ((0x000fffff & (2 + hx)) < 3) {           /* |f| < 2^-20 */
        if (f == zero) {
            if (k == 0) return zero;
            dk = (double)k; return dk * C_LN2_HI + dk * C_LN2_LO;
        }
        R = f * f * (0.5 - 0.33333333333333333 * f);
        if (k == 0) return f - R;
        dk = (double)k;
        return dk * C_LN2_HI - ((R - dk * C_LN2_LO) - f);
    }
    s = f / (2.0 + f);
    dk = (double)k;
    z = s * s;
    i = hx - 0x6147a;
    w = z * z;
    j = 0x6b851 - hx;
    t1 = w * (C_LG2 + w * (C_LG4 + w * C_LG6));
    t2 = z * (C_LG1 + w * (C_LG3 + w * (C_LG5 + w * C_LG7)));
    i |= j;
    R = t2 + t1;
    if (i > 0) {
        hfsq = 0.5 * f * f;
        if (k == 0) return f - (hfsq - s * (hfsq + R));
        return dk * C_LN2_HI - ((hfsq - (s * (hfsq + R) + dk * C_LN2_LO)) - f);
    }
    if (k == 0) return f - s * (f - R);
    return dk * C_LN2_HI - ((s * (f - R) - dk * C_LN2_LO) - f);
}

#define C_IVLN10     u2d(0x3FDBCB7B1526E50Eull)
#define C_LOG10_2HI  u2d(0x3FD34413509F6000ull)
#define C_LOG10_2LO  u2d(0x3D59FEF311F12B36ull)

double orc_log10(double x)
{
    const double zero = 0.0;
    int32_t hx = hi_word(x);
    uint32_t lx = lo_word(x);
    int32_t k = 0, i;
    double y, z;

    if (hx < 0x00100000) {
        if (((hx & 0x7fffffff) | lx) == 0) return -C_TWO54 / zero;
        if (hx < 0) return (x - x) / zero;
        k -= 54; x *= C_TWO54; hx = hi_word(x);
    }
    if (hx >= 0x7ff00000) return x + x;
    k += (hx >> 20) - 1023;
    i = (int32_t)(((uint32_t)k & 0x80000000u) >> 31);
    hx = (hx & 0x000fffff) | ((0x3ff - i) << 20);
    y = (double)(k + i);
    x = with_hi(x, hx);
    z = y * C_LOG10_2LO + C_IVLN10 * orc_log(x);
    return z + y * C_LOG10_2HI;
}

#define C_O_THRESH  u2d(0x40862E42FEFA39EFull)
#define C_U_THRESH  u2d(0xc0874910D52D3051ull)
#define C_INVLN2    u2d(0x3ff71547652b82feull)
#define C_P1        u2d(0x3FC555555555553Eull)
#define C_P2        u2d(0xBF66C16C16BEBD93ull)
#define C_P3        u2d(0x3F11566AAF25DE2Cull)
#define C_P4        u2d(0xBEBBBD41C5D26BF1ull)
#define C_P5        u2d(0x3E66376972BEA4D0ull)
#define C_TWOM1000  u2d(0x0170000000000000ull)

double orc_exp(double x)
{
    const double one = 1.0, huge = 1.0e+300;
    double y, hi = 0.0, lo = 0.0, c, t;
    int32_t k = 0, xsb;
    uint32_t hx;

    hx = (uint32_t)hi_word(x);
    xsb = (int32_t)((hx >> 31) & 1);
    hx &= 0x7fffffff;

    if (hx >= 0x40862E42) {                      /* |x| >= 709.78... */
        if (hx >= 0x7ff00000) {
            if (((hx & 0xfffff) | lo_word(x)) != 0) return x + x;   /* NaN */
            return (xsb == 0) ? x : 0.0;                            /* exp(+-inf) */
        }
        if (x > C_O_THRESH) return huge * huge;
        if (x < C_U_THRESH) return C_TWOM1000 * C_TWOM1000;
    }

    if (hx > 0x3fd62e42) {                       /* |x| > 0.5 ln2 */
        if (hx < 0x3FF0A2B2) {                   /* and |x| < 1.5 ln2 */
            if (xsb == 0) { hi = x - C_LN2_HI; lo = C_LN2_LO;  k = 1; }
            else          { hi = x + C_LN2_HI; lo = -C_LN2_LO; k = -1; }
        } else {
            k = (int32_t)(C_INVLN2 * x + (xsb == 0 ? 0.5 : -0.5));
            t = (double)k;
            hi = x - t * C_LN2_HI;               /* t*ln2HI is exact here */
            lo = t * C_LN2_LO;
        }
        x = hi - lo;
    } else if (hx < 0x3e300000) {                /* |x| < 2^-28 */
        if (huge + x > one) return one + x;
    } else {
        k = 0;
    }

    t = x * x;
    c = x - t * (C_P1 + t * (C_P2 + t * (C_P3 + t * (C_P4 + t * C_P5))));
    if (k == 0) return one - ((x * c) / (c - 2.0) - x);
    y = one - ((lo - (x * c) / (2.0 - c)) - hi);
    if (k >= -1021) {
        return with_hi(y, hi_word(y) + (k << 20));
    }
    y = with_hi(y, hi_word(y) + ((k + 1000) << 20));
    return y * C_TWOM1000;
}

#define C_LN10_OVER_10  u2d(0x3FCD791C5F888823ull)   /* 0.23025850929940458 = ln(10)/10 rounded */

double orc_dbm_to_mw(double dbm) { return orc_exp(dbm * C_LN10_OVER_10); }
double orc_mw_to_dbm(double mw)  { return 10.0 * orc_log10(mw); }
