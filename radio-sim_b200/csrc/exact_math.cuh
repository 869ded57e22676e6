// exact_math.cuh -- strict-IEEE double log / log10 / exp for the exact ("recheck") path.
//
// The SINR medium's decisions are DEFINED in double arithmetic with the classic fdlibm
// algorithms (the family java.lang.StrictMath is specified to use), evaluated without FMA
// contraction.  Every operation below is an explicit round-to-nearest intrinsic
// (__dadd_rn / __dmul_rn / __ddiv_rn), which nvcc never fuses, so the result is the same
// bit pattern any IEEE-754 machine produces for the same sequence.  tests/ compare these
// device functions bit-for-bit with the CPU oracle.
//
// The algorithms, evaluation order and coefficient tables follow fdlibm's e_log.c, e_log10.c and e_exp.c
// (third-party, not part of the reference), whose notice is preserved here as it requires:
//
// ====================================================
// Copyright (C) 1993 by Sun Microsystems, Inc. All rights reserved.
//
// Developed at SunSoft, a Sun Microsystems, Inc. business.
// Permission to use, copy, modify, and distribute this
// software is freely granted, provided that this notice
// is preserved.
// ====================================================
#pragma once
#include <cstdint>

namespace rmb {
namespace xm {

__device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double dvd(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ double bits(unsigned long long u) { return __longlong_as_double((long long)u); }
__device__ __forceinline__ double with_hi(double d, int hi) { return __hiloint2double(hi, __double2loint(d)); }

// natural logarithm, fdlibm e_log.c evaluation order
static __device__ __noinline__ double log_strict(double x)
{
    const double ln2_hi = bits(0x3fe62e42fee00000ull), ln2_lo = bits(0x3dea39ef35793c76ull);
    const double two54 = bits(0x4350000000000000ull);
    const double Lg1 = bits(0x3FE5555555555593ull), Lg2 = bits(0x3FD999999997FA04ull),
                 Lg3 = bits(0x3FD2492494229359ull), Lg4 = bits(0x3FCC71C51D8E78AFull),
                 Lg5 = bits(0x3FC7466496CB03DEull), Lg6 = bits(0x3FC39A09D078C69Full),
                 Lg7 = bits(0x3FC2F112DF3E5244ull);
    int hx = __double2hiint(x);
    unsigned lx = (unsigned)__double2loint(x);
    int k = 0;
    if (hx < 0x00100000) {
        if (((hx & 0x7fffffff) | lx) == 0) return dvd(-two54, 0.0);
        if (hx < 0) return dvd(sub(x, x), 0.0);
        k -= 54; x = mul(x, two54); hx = __double2hiint(x);
    }
    if (hx >= 0x7ff00000) return add(x, x);
    k += (hx >> 20) - 1023;
    hx &= 0x000fffff;
    int i = (hx + 0x95f64) & 0x100000;
    x = with_hi(x, hx | (i ^ 0x3ff00000));
    k += (i >> 20);
    double f = sub(x, 1.0);
    double dk = (double)k;
    if ((0x000fffff & (2 + hx)) < 3) {
        if (f == 0.0) {
            if (k == 0) return 0.0;
            return add(mul(dk, ln2_hi), mul(dk, ln2_lo));
        }
        double R = mul(mul(f, f), sub(0.5, mul(0.33333333333333333, f)));
        if (k == 0) return sub(f, R);
        return sub(mul(dk, ln2_hi), sub(sub(R, mul(dk, ln2_lo)), f));
    }
    double s = dvd(f, add(2.0, f));
    double z = mul(s, s);
    i = hx - 0x6147a;
    double w = mul(z, z);
    int j = 0x6b851 - hx;
    double t1 = mul(w, add(Lg2, mul(w, add(Lg4, mul(w, Lg6)))));
    double t2 = mul(z, add(Lg1, mul(w, add(Lg3, mul(w, add(Lg5, mul(w, Lg7)))))));
    i |= j;
    double R = add(t2, t1);
    if (i > 0) {
        double hfsq = mul(mul(0.5, f), f);
        if (k == 0) return sub(f, sub(hfsq, mul(s, add(hfsq, R))));
        return sub(mul(dk, ln2_hi), sub(sub(hfsq, add(mul(s, add(hfsq, R)), mul(dk, ln2_lo))), f));
    }
    if (k == 0) return sub(f, mul(s, sub(f, R)));
    return sub(mul(dk, ln2_hi), sub(sub(mul(s, sub(f, R)), mul(dk, ln2_lo)), f));
}

// fdlibm e_log10.c
__device__ __forceinline__ double log10_strict(double x)
{
    const double two54 = bits(0x4350000000000000ull);
    const double ivln10 = bits(0x3FDBCB7B1526E50Eull), log10_2hi = bits(0x3FD34413509F6000ull),
                 log10_2lo = bits(0x3D59FEF311F12B36ull);
    int hx = __double2hiint(x);
    unsigned lx = (unsigned)__double2loint(x);
    int k = 0;
    if (hx < 0x00100000) {
        if (((hx & 0x7fffffff) | lx) == 0) return dvd(-two54, 0.0);
        if (hx < 0) return dvd(sub(x, x), 0.0);
        k -= 54; x = mul(x, two54); hx = __double2hiint(x);
    }
    if (hx >= 0x7ff00000) return add(x, x);
    k += (hx >> 20) - 1023;
    int i = (int)(((unsigned)k & 0x80000000u) >> 31);
    hx = (hx & 0x000fffff) | ((0x3ff - i) << 20);
    double y = (double)(k + i);
    x = with_hi(x, hx);
    double z = add(mul(y, log10_2lo), mul(ivln10, log_strict(x)));
    return add(z, mul(y, log10_2hi));
}

// fdlibm e_exp.c
static __device__ __noinline__ double exp_strict(double x)
{
    const double ln2_hi = bits(0x3fe62e42fee00000ull), ln2_lo = bits(0x3dea39ef35793c76ull);
    const double o_thresh = bits(0x40862E42FEFA39EFull), u_thresh = bits(0xc0874910D52D3051ull);
    const double invln2 = bits(0x3ff71547652b82feull), twom1000 = bits(0x0170000000000000ull);
    const double P1 = bits(0x3FC555555555553Eull), P2 = bits(0xBF66C16C16BEBD93ull),
                 P3 = bits(0x3F11566AAF25DE2Cull), P4 = bits(0xBEBBBD41C5D26BF1ull),
                 P5 = bits(0x3E66376972BEA4D0ull);
    const double huge = 1.0e+300;
    double hi = 0.0, lo = 0.0;
    int k = 0;
    unsigned hx = (unsigned)__double2hiint(x);
    int xsb = (int)((hx >> 31) & 1);
    hx &= 0x7fffffff;
    if (hx >= 0x40862E42) {
        if (hx >= 0x7ff00000) {
            if (((hx & 0xfffff) | (unsigned)__double2loint(x)) != 0) return add(x, x);
            return (xsb == 0) ? x : 0.0;
        }
        if (x > o_thresh) return mul(huge, huge);
        if (x < u_thresh) return mul(twom1000, twom1000);
    }
    if (hx > 0x3fd62e42) {
        if (hx < 0x3FF0A2B2) {
            if (xsb == 0) { hi = sub(x, ln2_hi); lo = ln2_lo;  k = 1; }
            else          { hi = add(x, ln2_hi); lo = -ln2_lo; k = -1; }
        } else {
            k = __double2int_rz(add(mul(invln2, x), (xsb == 0 ? 0.5 : -0.5)));
            double t = (double)k;
            hi = sub(x, mul(t, ln2_hi));
            lo = mul(t, ln2_lo);
        }
        x = sub(hi, lo);
    } else if (hx < 0x3e300000) {
        if (add(huge, x) > 1.0) return add(1.0, x);
    } else {
        k = 0;
    }
    double t = mul(x, x);
    double c = sub(x, mul(t, add(P1, mul(t, add(P2, mul(t, add(P3, mul(t, add(P4, mul(t, P5))))))))));
    if (k == 0) return sub(1.0, sub(dvd(mul(x, c), sub(c, 2.0)), x));
    double y = sub(1.0, sub(sub(lo, dvd(mul(x, c), sub(2.0, c))), hi));
    if (k >= -1021) return with_hi(y, __double2hiint(y) + (k << 20));
    y = with_hi(y, __double2hiint(y) + ((k + 1000) << 20));
    return mul(y, twom1000);
}

// dBm <-> mW exactly as the SINR specification defines them
__device__ __forceinline__ double dbm_to_mw(double dbm) { return exp_strict(mul(dbm, bits(0x3FCD791C5F888823ull))); }
__device__ __forceinline__ double mw_to_dbm(double mw)  { return mul(10.0, log10_strict(mw)); }

// Position.getDistance (Position.java:56-64): unfused squares and adds, IEEE sqrt
__device__ __forceinline__ double distance(double x1, double y1, double z1, double x2, double y2, double z2)
{
    double dx = sub(x1, x2), dy = sub(y1, y2), dz = sub(z1, z2);
    dx = mul(dx, dx); dy = mul(dy, dy); dz = mul(dz, dz);
    return __dsqrt_rn(add(add(dx, dy), dz));
}

}  // namespace xm
}  // namespace rmb
