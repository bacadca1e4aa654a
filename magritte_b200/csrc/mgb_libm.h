// mgb_libm.h -- bit-exact single-precision libm for the device.
//
// The reference evaluates five SINGLE-precision libm functions on double arguments inside its fp64
// hot path: expf (Gaussian profile, reference src/solver/solver.tpp:1109 and
// src/model/thermodynamics/thermodynamics.tpp:13), expm1f (Planck function, solver.tpp:1120),
// erff (large-shift optical depth, solver.tpp:2450), logf and powf (sub-step count and log
// interpolation, src/solver/solver_interp_helper.tpp:37,39,67,68,98). A 1e-10 match of J therefore
// needs the SAME float results as the host libm the reference links against, bit for bit.
//
// That libm is a third-party dependency absent from /root/reference: GNU libc 2.39 (x86-64). Its
// algorithms are restated here from the published sources/papers:
//   * expf, logf, powf: the ARM "optimized-routines" single-precision kernels adopted by glibc 2.28
//     (table + short polynomial evaluated in double, one final rounding to float). On x86-64 the
//     multiarch build selects the FMA-compiled variant at run time (ifunc); the placement of fused
//     multiply-adds below is the one in that variant, so every fma() here is deliberate.
//   * erff, expm1f: the Sun fdlibm float ports (s_erff.c, s_expm1f.c), plain float arithmetic with
//     no fused operations; erff calls expf for |x| >= 0.84375.
// Table entries and polynomial coefficients are those of glibc 2.39's __exp2f_data, __logf_data and
// __powf_log2_data objects and of fdlibm.
//
// Proof of bit-exactness: tests/test_libm_host.py compiles this header for the host (explicit fma(),
// -ffp-contract=off) and compares against the box's own libm over ALL 2^32 float inputs (expf, logf,
// erff, expm1f) and a 2^32-point lattice for powf; on the GPU, mgb_selftest_libm() runs the device
// build against host-libm results (tests/test_gpu_parity.py::test_device_libm_bit_exact).
//
// Notices of the restated third-party sources (algorithms, tables and coefficients; the code below is this project's):
//   * expf / logf / powf kernels and their data tables: Arm Optimized Routines, Copyright (c) 2017-2018, Arm Limited,
//     SPDX-License-Identifier: MIT (as imported into glibc 2.28+ under the GNU LGPL v2.1 or later, sysdeps/ieee754/flt-32/e_expf.c,
//     e_logf.c, e_powf.c, e_exp2f_data.c, e_logf_data.c, e_powf_log2_data.c).
//   * erff / expm1f: fdlibm s_erf.c / s_expm1.c float ports. "Copyright (C) 1993 by Sun Microsystems, Inc. All rights reserved.
//     Developed at SunPro, a Sun Microsystems, Inc. business. Permission to use, copy, modify, and distribute this software is
//     freely granted, provided that this notice is preserved."
//
// The translation unit that includes this must be compiled with nvcc -fmad=false (no implicit
// contraction); -prec-div / -prec-sqrt / no -ftz are the nvcc defaults.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define MGB_LIBM_FN static __device__ __forceinline__
#define MGB_LIBM_TAB static __device__ const
#define MGB_FMA(a, b, c) __fma_rn((a), (b), (c))
#define MGB_F2I(x) __float_as_int(x)
#define MGB_I2F(i) __int_as_float(i)
#define MGB_D2LL(x) __double_as_longlong(x)
#define MGB_LL2D(i) __longlong_as_double(i)
#else
#include <math.h>
#include <string.h>
#define MGB_LIBM_FN static inline
#define MGB_LIBM_TAB static const
#define MGB_FMA(a, b, c) fma((a), (b), (c))
static inline int32_t MGB_F2I(float x) { int32_t i; memcpy(&i, &x, 4); return i; }
static inline float MGB_I2F(int32_t i) { float x; memcpy(&x, &i, 4); return x; }
static inline int64_t MGB_D2LL(double x) { int64_t i; memcpy(&i, &x, 8); return i; }
static inline double MGB_LL2D(int64_t i) { double x; memcpy(&x, &i, 8); return x; }
#endif

// ---- __exp2f_data: T[i] = bits(2^(i/32)) - (i << 47), i = 0..31 (tools/gen_exp2f_table.py regenerates it)
MGB_LIBM_TAB uint64_t mgb_exp2f_tab[32] = {
    0x3ff0000000000000ull, 0x3fefd9b0d3158574ull, 0x3fefb5586cf9890full, 0x3fef9301d0125b51ull,
    0x3fef72b83c7d517bull, 0x3fef54873168b9aaull, 0x3fef387a6e756238ull, 0x3fef1e9df51fdee1ull,
    0x3fef06fe0a31b715ull, 0x3feef1a7373aa9cbull, 0x3feedea64c123422ull, 0x3feece086061892dull,
    0x3feebfdad5362a27ull, 0x3feeb42b569d4f82ull, 0x3feeab07dd485429ull, 0x3feea47eb03a5585ull,
    0x3feea09e667f3bcdull, 0x3fee9f75e8ec5f74ull, 0x3feea11473eb0187ull, 0x3feea589994cce13ull,
    0x3feeace5422aa0dbull, 0x3feeb737b0cdc5e5ull, 0x3feec49182a3f090ull, 0x3feed503b23e255dull,
    0x3feee89f995ad3adull, 0x3feeff76f2fb5e47ull, 0x3fef199bdd85529cull, 0x3fef3720dcef9069ull,
    0x3fef5818dcfba487ull, 0x3fef7c97337b9b5full, 0x3fefa4afa2a490daull, 0x3fefd0765b6e4540ull};

#ifndef MGB_EXP2F_TAB /* where the 2^(i/32) table is read from (a device build may stage it in shared memory) */
#define MGB_EXP2F_TAB(i) mgb_exp2f_tab[i]
#endif
// expf used inside erff: device code may route it to a shared out-of-line instance (mgb_device.cuh)
#ifndef MGB_ERFF_EXPF
#define MGB_ERFF_EXPF mgb_expf
#endif

// ---- __logf_data / __powf_log2_data share invc; logc is ln(c) resp. log2(c)
MGB_LIBM_TAB double mgb_log_invc[16] = {
    0x1.661ec79f8f3bep+0, 0x1.571ed4aaf883dp+0, 0x1.49539f0f010b0p+0, 0x1.3c995b0b80385p+0,
    0x1.30d190c8864a5p+0, 0x1.25e227b0b8ea0p+0, 0x1.1bb4a4a1a343fp+0, 0x1.12358f08ae5bap+0,
    0x1.0953f419900a7p+0, 0x1.0000000000000p+0, 0x1.e608cfd9a47acp-1, 0x1.ca4b31f026aa0p-1,
    0x1.b2036576afce6p-1, 0x1.9c2d163a1aa2dp-1, 0x1.886e6037841edp-1, 0x1.767dcf5534862p-1};
MGB_LIBM_TAB double mgb_logf_logc[16] = {
    -0x1.57bf7808caadep-2, -0x1.2bef0a7c06ddbp-2, -0x1.01eae7f513a67p-2, -0x1.b31d8a68224e9p-3,
    -0x1.6574f0ac07758p-3, -0x1.1aa2bc79c8100p-3, -0x1.a4e76ce8c0e5ep-4, -0x1.1973c5a611cccp-4,
    -0x1.252f438e10c1ep-5, 0x0.0p+0,              0x1.aa5aa5df25984p-5,  0x1.c5e53aa362eb4p-4,
    0x1.526e57720db08p-3,  0x1.bc2860d224770p-3,  0x1.1058bc8a07ee1p-2,  0x1.4043057b6ee09p-2};
MGB_LIBM_TAB double mgb_powf_log2c[16] = {
    -0x1.efec65b963019p-2, -0x1.b0b6832d4fca4p-2, -0x1.7418b0a1fb77bp-2, -0x1.39de91a6dcf7bp-2,
    -0x1.01d9bf3f2b631p-2, -0x1.97c1d1b3b7af0p-3, -0x1.2f9e393af3c9fp-3, -0x1.960cbbf788d5cp-4,
    -0x1.a6f9db6475fcep-5, 0x0.0p+0,              0x1.338ca9f24f53dp-4,  0x1.476a9543891bap-3,
    0x1.e840b4ac4e4d2p-3,  0x1.40645f0c6651cp-2,  0x1.88e9c2c1b9ff8p-2,  0x1.ce0a44eb17bccp-2};

// 2^(k/32) * (1 + C2' r + C1' r^2 + C0' r^3) core shared by expf (coefficients pre-scaled by 1/32^i)
// and powf (unscaled); `ki` carries k in its low bits (the "shift" trick), sign_bias goes to bit 63.
MGB_LIBM_FN float mgb_exp2f_core(double r, uint64_t ki, uint64_t sign_bias, double C0, double C1, double C2) {
    uint64_t t = MGB_EXP2F_TAB(ki & 31);
    t += (ki + sign_bias) << 47;
    const double s  = MGB_LL2D((int64_t)t);
    const double z  = MGB_FMA(r, C0, C1);
    const double r2 = r * r;
    double y        = MGB_FMA(r, C2, 1.0);
    y               = MGB_FMA(z, r2, y);
    y               = y * s;
    return (float)y;
}

// glibc __expf (FMA variant). reference call sites: solver.tpp:1109, thermodynamics.tpp:13.
MGB_LIBM_FN float mgb_expf(float x) {
    const uint32_t ux     = (uint32_t)MGB_F2I(x);
    const uint32_t abstop = (ux >> 20) & 0x7ff;
    if (abstop >= 0x42b) { /* |x| >= 88 or NaN */
        if (ux == 0xff800000u) return 0.0f;
        if (abstop >= 0x7f8) return x + x;
        if (x > 0x1.62e42ep6f) return MGB_I2F(0x7f800000);  /* overflow -> +inf (round to nearest) */
        if (x < -0x1.9fe368p6f) return 0.0f;                /* underflow -> +0 */
        if (x < -0x1.9d1d9ep6f) return 0x1.4p-75f * 0x1.4p-75f; /* may-underflow value: 2^-149 */
    }
    const double xd    = (double)x;
    const double SHIFT = 0x1.8p+52, InvLn2N = 0x1.71547652b82fep+5;
    double kd          = MGB_FMA(InvLn2N, xd, SHIFT);
    const uint64_t ki  = (uint64_t)MGB_D2LL(kd);
    kd                 = kd - SHIFT;
    const double r     = MGB_FMA(InvLn2N, xd, -kd);
    return mgb_exp2f_core(r, ki, 0, 0x1.c6af84b912394p-20, 0x1.ebfce50fac4f3p-13, 0x1.62e42ff0c52d6p-6);
}

// glibc __logf (FMA variant). reference call sites: solver_interp_helper.tpp:37,39,67,68.
MGB_LIBM_FN float mgb_logf(float x) {
    uint32_t ix = (uint32_t)MGB_F2I(x);
    if (ix == 0x3f800000u) return 0.0f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2 == 0) return MGB_I2F((int32_t)0xff800000);          /* log(+-0) = -inf */
        if (ix == 0x7f800000u) return x;                               /* log(inf) = inf */
        if ((ix & 0x80000000u) || ix * 2 >= 0xff000000u) return (x - x) / (x - x); /* invalid -> NaN */
        ix = (uint32_t)MGB_F2I(x * 0x1p23f); /* subnormal: normalise */
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i        = (int)((tmp >> 19) & 15);
    const int k        = (int32_t)tmp >> 23;
    const uint32_t iz  = ix - (tmp & 0xff800000u);
    const double invc = mgb_log_invc[i], logc = mgb_logf_logc[i];
    const double z  = (double)MGB_I2F((int32_t)iz);
    const double r  = MGB_FMA(z, invc, -1.0);
    const double y0 = MGB_FMA((double)k, 0x1.62e42fefa39efp-1, logc);
    const double r2 = r * r;
    double y        = MGB_FMA(r, 0x1.5575b0be00b6ap-2, -0x1.ffffef20a4123p-2);
    y               = MGB_FMA(r2, -0x1.00ea348b88334p-2, y);
    y               = MGB_FMA(r2, y, r + y0);
    return (float)y;
}

MGB_LIBM_FN int mgb_powf_checkint(uint32_t iy) { /* 0 non-integer, 1 odd, 2 even */
    const int e = (int)((iy >> 23) & 0xff);
    if (e < 0x7f) return 0;
    if (e > 0x7f + 23) return 2;
    if (iy & ((1u << (0x7f + 23 - e)) - 1)) return 0;
    if (iy & (1u << (0x7f + 23 - e))) return 1;
    return 2;
}
MGB_LIBM_FN int mgb_zeroinfnan(uint32_t ix) { return 2 * ix - 1 >= 2u * 0x7f800000u - 1; }

// glibc __powf (FMA variant). reference call site: solver_interp_helper.tpp:98.
MGB_LIBM_FN float mgb_powf(float x, float y) {
    uint32_t sign_bias = 0;
    uint32_t ix = (uint32_t)MGB_F2I(x);
    const uint32_t iy = (uint32_t)MGB_F2I(y);
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u || mgb_zeroinfnan(iy)) {
        if (mgb_zeroinfnan(iy)) {
            if (2 * iy == 0) return 1.0f;
            if (ix == 0x3f800000u) return 1.0f;
            if (2 * ix > 2u * 0x7f800000u || 2 * iy > 2u * 0x7f800000u) return x + y;
            if (2 * ix == 2 * 0x3f800000u) return 1.0f;
            if ((2 * ix < 2 * 0x3f800000u) == !(iy & 0x80000000u)) return 0.0f;
            return y * y;
        }
        if (mgb_zeroinfnan(ix)) {
            float x2 = x * x;
            if ((ix & 0x80000000u) && mgb_powf_checkint(iy) == 1) { x2 = -x2; sign_bias = 1; }
            if (2 * ix == 0 && (iy & 0x80000000u)) return sign_bias ? MGB_I2F((int32_t)0xff800000) : MGB_I2F(0x7f800000);
            return (iy & 0x80000000u) ? 1 / x2 : x2;
        }
        if (ix & 0x80000000u) {
            const int yint = mgb_powf_checkint(iy);
            if (yint == 0) return (x - x) / (x - x);
            if (yint == 1) sign_bias = 1u << 16; /* SIGN_BIAS = 1 << (EXP2F_TABLE_BITS + 11) */
            ix &= 0x7fffffffu;
        }
        if (ix < 0x00800000u) {
            ix = (uint32_t)MGB_F2I(x * 0x1p23f);
            ix &= 0x7fffffffu;
            ix -= 23u << 23;
        }
    }
    /* log2_inline */
    const uint32_t tmp = ix - 0x3f330000u;
    const int i        = (int)((tmp >> 19) & 15);
    const uint32_t top = tmp & 0xff800000u;
    const uint32_t iz  = ix - top;
    const int k        = (int32_t)top >> 23;
    const double invc = mgb_log_invc[i], logc = mgb_powf_log2c[i];
    const double z  = (double)MGB_I2F((int32_t)iz);
    const double r  = MGB_FMA(z, invc, -1.0);
    const double y0 = logc + (double)k;
    double yy       = MGB_FMA(r, 0x1.27616c9496e0bp-2, -0x1.71969a075c67ap-2);
    const double p  = MGB_FMA(r, 0x1.ec70a6ca7baddp-2, -0x1.7154748bef6c8p-1);
    const double r2 = r * r;
    double q        = MGB_FMA(r, 0x1.71547652ab82bp+0, y0);
    const double r4 = r2 * r2;
    q               = MGB_FMA(r2, p, q);
    yy              = MGB_FMA(yy, r4, q);
    const double ylogx = (double)y * yy;
    if ((((uint64_t)MGB_D2LL(ylogx) >> 47) & 0xffff) >= (0x405f800000000000ull >> 47)) { /* |y log2 x| >= 126 */
        if (ylogx > 0x1.fffffffd1d571p+6) return sign_bias ? MGB_I2F((int32_t)0xff800000) : MGB_I2F(0x7f800000);
        if (ylogx <= -150.0) return sign_bias ? -0.0f : 0.0f;
        if (ylogx < -149.0) { const float v = 0x1.4p-75f * 0x1.4p-75f; return sign_bias ? -v : v; }
    }
    /* exp2_inline */
    const double SHIFT = 0x1.8p+47;
    double kd          = ylogx + SHIFT;
    const uint64_t ki  = (uint64_t)MGB_D2LL(kd);
    kd                 = kd - SHIFT;
    const double rr    = ylogx - kd;
    return mgb_exp2f_core(rr, ki, sign_bias, 0x1.c6af84b912394p-5, 0x1.ebfce50fac4f3p-3, 0x1.62e42ff0c52d6p-1);
}

// fdlibm s_expm1f.c (glibc 2.39 generic build: float arithmetic, no fused operations).
// reference call site: solver.tpp:1120 (Planck function).
MGB_LIBM_FN float mgb_expm1f(float x) {
    const float one = 1.0f, huge = 1.0e+30f, tiny = 1.0e-30f, o_threshold = 8.8721679688e+01f,
                ln2_hi = 6.9313812256e-01f, ln2_lo = 9.0580006145e-06f, invln2 = 1.4426950216e+00f,
                Q1 = -3.3333335072e-02f, Q2 = 1.5873016091e-03f, Q3 = -7.9365076090e-05f, Q4 = 4.0082177293e-06f,
                Q5 = -2.0109921195e-07f;
    float y, hi, lo, c = 0.0f, t, e, hxs, hfx, r1;
    int32_t k;
    uint32_t hx        = (uint32_t)MGB_F2I(x);
    const uint32_t xsb = hx & 0x80000000u;
    hx &= 0x7fffffffu;
    if (hx >= 0x4195b844u) { /* |x| >= 27 ln2 */
        if (hx >= 0x42b17218u) {
            if (hx > 0x7f800000u) return x + x;
            if (hx == 0x7f800000u) return (xsb == 0) ? x : -1.0f;
            if (x > o_threshold) return huge * huge;
        }
        if (xsb != 0) return tiny - one;
    }
    if (hx > 0x3eb17218u) { /* |x| > 0.5 ln2 */
        if (hx < 0x3F851592u) {
            if (xsb == 0) { hi = x - ln2_hi; lo = ln2_lo; k = 1; }
            else { hi = x + ln2_hi; lo = -ln2_lo; k = -1; }
        } else {
            k  = (int32_t)(invln2 * x + ((xsb == 0) ? 0.5f : -0.5f));
            t  = (float)k;
            hi = x - t * ln2_hi;
            lo = t * ln2_lo;
        }
        x = hi - lo;
        c = (hi - x) - lo;
    } else if (hx < 0x33000000u) {
        return x;
    } else
        k = 0;
    hfx = 0.5f * x;
    hxs = x * hfx;
    r1  = one + hxs * (Q1 + hxs * (Q2 + hxs * (Q3 + hxs * (Q4 + hxs * Q5))));
    t   = 3.0f - r1 * hfx;
    e   = hxs * ((r1 - t) / (6.0f - x * t));
    if (k == 0) return x - (x * e - hxs);
    e = (x * (e - c) - c);
    e -= hxs;
    if (k == -1) return 0.5f * (x - e) - 0.5f;
    if (k == 1) {
        if (x < -0.25f) return -2.0f * (e - (x + 0.5f));
        return one + 2.0f * (x - e);
    }
    if (k <= -2 || k > 56) {
        y = one - (e - x);
        y = MGB_I2F(MGB_F2I(y) + (int32_t)((uint32_t)k << 23));
        return y - one;
    }
    if (k < 23) {
        t = MGB_I2F(0x3f800000 - (0x1000000 >> k));
        y = t - (e - x);
        y = MGB_I2F(MGB_F2I(y) + (k << 23));
    } else {
        t = MGB_I2F((0x7f - k) << 23);
        y = x - (e + t);
        y += one;
        y = MGB_I2F(MGB_F2I(y) + (k << 23));
    }
    return y;
}

// fdlibm s_erff.c (glibc 2.39 generic build). reference call site: solver.tpp:2450.
MGB_LIBM_FN float mgb_erff(float x) {
    const float one = 1.0f, tiny = 1e-30f, erx = 8.4506291151e-01f, efx = 1.2837916613e-01f,
    pp0 = 1.2837916613e-01f, pp1 = -3.2504209876e-01f, pp2 = -2.8481749818e-02f, pp3 = -5.7702702470e-03f,
    pp4 = -2.3763017452e-05f, qq1 = 3.9791721106e-01f, qq2 = 6.5022252500e-02f, qq3 = 5.0813062117e-03f,
    qq4 = 1.3249473704e-04f, qq5 = -3.9602282413e-06f,
    pa0 = -2.3621185683e-03f, pa1 = 4.1485610604e-01f, pa2 = -3.7220788002e-01f, pa3 = 3.1834661961e-01f,
    pa4 = -1.1089469492e-01f, pa5 = 3.5478305072e-02f, pa6 = -2.1663755178e-03f, qa1 = 1.0642088205e-01f,
    qa2 = 5.4039794207e-01f, qa3 = 7.1828655899e-02f, qa4 = 1.2617121637e-01f, qa5 = 1.3637083583e-02f,
    qa6 = 1.1984500103e-02f,
    ra0 = -9.8649440333e-03f, ra1 = -6.9385856390e-01f, ra2 = -1.0558626175e+01f, ra3 = -6.2375331879e+01f,
    ra4 = -1.6239666748e+02f, ra5 = -1.8460508728e+02f, ra6 = -8.1287437439e+01f, ra7 = -9.8143291473e+00f,
    sa1 = 1.9651271820e+01f, sa2 = 1.3765776062e+02f, sa3 = 4.3456588745e+02f, sa4 = 6.4538726807e+02f,
    sa5 = 4.2900814819e+02f, sa6 = 1.0863500214e+02f, sa7 = 6.5702495575e+00f, sa8 = -6.0424413532e-02f,
    rb0 = -9.8649431020e-03f, rb1 = -7.9928326607e-01f, rb2 = -1.7757955551e+01f, rb3 = -1.6063638306e+02f,
    rb4 = -6.3756646729e+02f, rb5 = -1.0250950928e+03f, rb6 = -4.8351919556e+02f, sb1 = 3.0338060379e+01f,
    sb2 = 3.2579251099e+02f, sb3 = 1.5367296143e+03f, sb4 = 3.1998581543e+03f, sb5 = 2.5530502930e+03f,
    sb6 = 4.7452853394e+02f, sb7 = -2.2440952301e+01f;
    float R, S, P, Q, s, y, z, r;
    const int32_t hx = MGB_F2I(x);
    const int32_t ix = hx & 0x7fffffff;
    if (ix >= 0x7f800000) {
        const int32_t i = (int32_t)(((uint32_t)hx >> 31) << 1);
        return (float)(1 - i) + one / x;
    }
    if (ix < 0x3f580000) { /* |x| < 0.84375 */
        if (ix < 0x31800000) {
            if (ix < 0x04000000) return 0.0625f * (16.0f * x + (16.0f * efx) * x);
            return x + efx * x;
        }
        z = x * x;
        r = pp0 + z * (pp1 + z * (pp2 + z * (pp3 + z * pp4)));
        s = one + z * (qq1 + z * (qq2 + z * (qq3 + z * (qq4 + z * qq5))));
        y = r / s;
        return x + x * y;
    }
    if (ix < 0x3fa00000) { /* 0.84375 <= |x| < 1.25 */
        s = MGB_I2F(ix) - one;
        P = pa0 + s * (pa1 + s * (pa2 + s * (pa3 + s * (pa4 + s * (pa5 + s * pa6)))));
        Q = one + s * (qa1 + s * (qa2 + s * (qa3 + s * (qa4 + s * (qa5 + s * qa6)))));
        if (hx >= 0) return erx + P / Q;
        return -erx - P / Q;
    }
    if (ix >= 0x40c00000) { /* |x| >= 6 */
        if (hx >= 0) return one - tiny;
        return tiny - one;
    }
    x = MGB_I2F(ix);
    s = one / (x * x);
    if (ix < 0x4036DB6E) { /* |x| < 1/0.35 */
        R = ra0 + s * (ra1 + s * (ra2 + s * (ra3 + s * (ra4 + s * (ra5 + s * (ra6 + s * ra7))))));
        S = one + s * (sa1 + s * (sa2 + s * (sa3 + s * (sa4 + s * (sa5 + s * (sa6 + s * (sa7 + s * sa8)))))));
    } else {
        R = rb0 + s * (rb1 + s * (rb2 + s * (rb3 + s * (rb4 + s * (rb5 + s * rb6)))));
        S = one + s * (sb1 + s * (sb2 + s * (sb3 + s * (sb4 + s * (sb5 + s * (sb6 + s * sb7))))));
    }
    z = MGB_I2F(ix & (int32_t)0xfffff000);
    r = MGB_ERFF_EXPF(-z * z - 0.5625f) * MGB_ERFF_EXPF((z - x) * (z + x) + R / S);
    if (hx >= 0) return one - r / x;
    return r / x - one;
}

// ---------------------------------------------------------------------------------------------------
// erff without the range branches (the solve kernel's large-shift path evaluates two erff per segment and bin, and the lanes of a
// warp fall into different |x| ranges: the branchy routine above ran at 12 of 32 lanes). Same arithmetic, same bits:
//   * the four rational approximations share one Horner evaluation on a per-range coefficient row, rows padded with LEADING zeros
//     (c + t * 0 = c exactly for finite t, so the padded steps reproduce the shorter polynomial bit for bit);
//   * the variable is x*x, |x| - 1 or 1 / (x*x) by range; the exp() factor of the two outer ranges is evaluated by every lane on its
//     own (possibly meaningless) argument with the special-case-free expf kernel and selected at the end;
//   * |x| >= 6 gives +-(1 - tiny) = +-1; inf, NaN and |x| < 2^-28 go to mgb_erff.
// tests/test_libm_host.py compares it with the box's erff over all 2^32 inputs.
MGB_LIBM_TAB float mgb_erff_rows[4][16] = {
    /* |x| < 0.84375: pp0..pp4, qq1..qq5 */
    {1.2837916613e-01f, -3.2504209876e-01f, -2.8481749818e-02f, -5.7702702470e-03f, -2.3763017452e-05f, 0.0f, 0.0f, 0.0f,
     3.9791721106e-01f, 6.5022252500e-02f, 5.0813062117e-03f, 1.3249473704e-04f, -3.9602282413e-06f, 0.0f, 0.0f, 0.0f},
    /* 0.84375 <= |x| < 1.25: pa0..pa6, qa1..qa6 */
    {-2.3621185683e-03f, 4.1485610604e-01f, -3.7220788002e-01f, 3.1834661961e-01f, -1.1089469492e-01f, 3.5478305072e-02f,
     -2.1663755178e-03f, 0.0f, 1.0642088205e-01f, 5.4039794207e-01f, 7.1828655899e-02f, 1.2617121637e-01f, 1.3637083583e-02f,
     1.1984500103e-02f, 0.0f, 0.0f},
    /* 1.25 <= |x| < 1/0.35: ra0..ra7, sa1..sa8 */
    {-9.8649440333e-03f, -6.9385856390e-01f, -1.0558626175e+01f, -6.2375331879e+01f, -1.6239666748e+02f, -1.8460508728e+02f,
     -8.1287437439e+01f, -9.8143291473e+00f, 1.9651271820e+01f, 1.3765776062e+02f, 4.3456588745e+02f, 6.4538726807e+02f,
     4.2900814819e+02f, 1.0863500214e+02f, 6.5702495575e+00f, -6.0424413532e-02f},
    /* 1/0.35 <= |x| < 6: rb0..rb6, sb1..sb7 */
    {-9.8649431020e-03f, -7.9928326607e-01f, -1.7757955551e+01f, -1.6063638306e+02f, -6.3756646729e+02f, -1.0250950928e+03f,
     -4.8351919556e+02f, 0.0f, 3.0338060379e+01f, 3.2579251099e+02f, 1.5367296143e+03f, 3.1998581543e+03f, 2.5530502930e+03f,
     4.7452853394e+02f, -2.2440952301e+01f, 0.0f}};
#ifndef MGB_ERFF_ROW /* where a coefficient row is read from (a device build may stage the table in shared memory) */
#define MGB_ERFF_ROW(r) mgb_erff_rows[r]
#endif

// __expf's table + polynomial kernel without its special cases: the bits of mgb_expf(x) for |x| < 88; any other input gives some
// float without trapping (the callers discard it).
MGB_LIBM_FN float mgb_expf_kernel(float x) {
    const double xd    = (double)x;
    const double SHIFT = 0x1.8p+52, InvLn2N = 0x1.71547652b82fep+5;
    double kd          = MGB_FMA(InvLn2N, xd, SHIFT);
    const uint64_t ki  = (uint64_t)MGB_D2LL(kd);
    kd                 = kd - SHIFT;
    const double r     = MGB_FMA(InvLn2N, xd, -kd);
    return mgb_exp2f_core(r, ki, 0, 0x1.c6af84b912394p-20, 0x1.ebfce50fac4f3p-13, 0x1.62e42ff0c52d6p-6);
}

// The straight-line part: valid for 2^-28 <= |x| < inf (the caller routes the rest to mgb_erff).
MGB_LIBM_FN float mgb_erff_main(float x) {
    const float erx  = 8.4506291151e-01f;
    const int32_t hx = MGB_F2I(x), ix = hx & 0x7fffffff;
    const int far_   = ix >= 0x40c00000; /* |x| >= 6: the result is +-(1 - tiny); the arithmetic below runs on a harmless stand-in */
    const float ax   = far_ ? 3.0f : MGB_I2F(ix);
    const int r      = (ix < 0x3f580000) ? 0 : (ix < 0x3fa00000) ? 1 : (ix < 0x4036DB6E) ? 2 : 3;
    const float xx   = ax * ax;
    const float t    = (r == 0) ? xx : (r == 1) ? ax - 1.0f : 1.0f / xx;
    const float* c   = MGB_ERFF_ROW(r);
    float P = c[7], Q = c[15];
    P = c[6] + t * P; P = c[5] + t * P; P = c[4] + t * P; P = c[3] + t * P; P = c[2] + t * P; P = c[1] + t * P; P = c[0] + t * P;
    Q = c[14] + t * Q; Q = c[13] + t * Q; Q = c[12] + t * Q; Q = c[11] + t * Q; Q = c[10] + t * Q; Q = c[9] + t * Q; Q = c[8] + t * Q;
    Q = 1.0f + t * Q;
    const float y  = P / Q;
    const float z  = MGB_I2F(MGB_F2I(ax) & (int32_t)0xfffff000);
    const float e1 = mgb_expf_kernel(-z * z - 0.5625f), e2 = mgb_expf_kernel((z - ax) * (z + ax) + y);
    const float rr = e1 * e2;
    float big      = 1.0f - rr / ax;
    big            = far_ ? 1.0f : big; /* one - tiny */
    const float mag = (r == 1) ? erx + y : big;
    return (r == 0) ? x + x * y : ((hx < 0) ? -mag : mag);
}

MGB_LIBM_FN float mgb_erff_u(float x) {
    const int32_t ix = MGB_F2I(x) & 0x7fffffff;
    if (ix >= 0x7f800000 || ix < 0x31800000) return mgb_erff(x);
    return mgb_erff_main(x);
}
