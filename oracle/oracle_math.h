/*
 * oracle_math.h -- TEST INFRASTRUCTURE ONLY (see oracle/README.md).
 *
 * Deterministic double-precision log / log10 / exp used by the SINR medium
 * specification.  The algorithms are the classic fdlibm ones (the same family
 * Java's StrictMath is specified to use), written out here with nothing but
 * IEEE-754 +,-,*,/ and integer bit manipulation, so that any compiler that
 * does NOT contract a*b+c into an FMA produces bit-identical results.
 * Compile with -ffp-contract=off.
 *
 * The product's device code (radio-sim_b200/csrc/exact_math.cuh) implements
 * the same algorithms independently with __dmul_rn/__dadd_rn; tests compare
 * the two bit-for-bit.
 */
#ifndef ORACLE_MATH_H
#define ORACLE_MATH_H

#ifdef __cplusplus
extern "C" {
#endif

double orc_log(double x);     /* natural log, fdlibm e_log algorithm        */
double orc_log10(double x);   /* fdlibm e_log10 algorithm (uses orc_log)    */
double orc_exp(double x);     /* fdlibm e_exp algorithm                     */

/* dBm <-> mW as the SINR spec defines them (DESIGN.md "exact arithmetic") */
double orc_dbm_to_mw(double dbm);   /* orc_exp(dbm * (ln10/10))   */
double orc_mw_to_dbm(double mw);    /* 10.0 * orc_log10(mw)       */

#ifdef __cplusplus
}
#endif
#endif
