// sinf / cosf with the exact arithmetic of glibc 2.39's x86-64 FMA variants (__sinf_fma, __cosf_fma,
// __sincosf_fma: sysdeps/ieee754/flt-32/s_sinf.c, s_cosf.c, sincosf.h — the ARM optimized-routines
// algorithm). The reference's CPU build calls these through libm for every quaternion it builds
// (Quat::angleAxis) and every box it orients (obb.hpp); restating them operation for operation
// (same double constants, same fused multiply-adds, one final rounding to float) makes the CUDA
// per-agent state bit-identical to the CPU reference instead of "within an ulp", so no error can
// accumulate through the yaw -> quaternion -> yaw round trip of every step.
//
// Published algorithm: reduce x to r = x - n*(pi/2) with n = round(x * 2/pi) (done as a scaled
// float->int conversion), then evaluate a degree-7 sine or degree-8 cosine polynomial in double.
// The instruction order below was read from the disassembly of this image's libm.so.6 and is
// pinned by tests/test_cpu_trig.py (bit-equality with libm over 2^24+ inputs).
// Valid for |x| < 120; callers fall back to fp64 sin/cos beyond that (never hit: |yaw| <= 2*pi).
//
// Attribution: the algorithm and the coefficient tables are glibc's (GNU C Library 2.39, LGPL-2.1-or-later,
// Copyright (C) Free Software Foundation, Inc.), which took them from ARM's optimized-routines
// (Copyright (c) Arm Limited, MIT OR Apache-2.0 WITH LLVM-exception). Third-party arithmetic restated for
// bit-parity; it is not code of the reference repository.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#if defined(__CUDA_ARCH__)
#define GDT_FN __device__ __forceinline__
#define GDT_MUL(a, b) __dmul_rn((a), (b))
#define GDT_FMA(a, b, c) __fma_rn((a), (b), (c))
#define GDT_CONST static const __device__
#else
#define GDT_FN static inline
#define GDT_MUL(a, b) ((a) * (b))
#define GDT_FMA(a, b, c) __builtin_fma((a), (b), (c))
#endif

namespace gdtrig {

// __sincosf_table[0] / [1] (the cosine coefficients are negated in the second entry)
struct Tab {
    double c0, c1, s1, c2, s2, c3, s3, c4;
};
GDT_FN Tab tab(int neg) {
    Tab t;
    const double sg = neg ? -1.0 : 1.0;
    t.c0 = sg * 0x1p0;
    t.c1 = sg * -0x1.ffffffd0c621cp-2;
    t.s1 = -0x1.555545995a603p-3;
    t.c2 = sg * 0x1.55553e1068f19p-5;
    t.s2 = 0x1.1107605230bc4p-7;
    t.c3 = sg * -0x1.6c087e89a359dp-10;
    t.s3 = -0x1.994eb3774cf24p-13;
    t.c4 = sg * 0x1.99343027bf8c3p-16;
    return t;
}

GDT_FN double poly_sin(double x, double x2, const Tab &p) {
    const double s1 = GDT_FMA(x2, p.s3, p.s2);
    const double x3 = GDT_MUL(x2, x);
    const double x7 = GDT_MUL(x2, x3);
    const double s = GDT_FMA(x3, p.s1, x);
    return GDT_FMA(s1, x7, s);
}

GDT_FN double poly_cos(double x2, const Tab &p) {
    const double x4 = GDT_MUL(x2, x2);
    const double c1 = GDT_FMA(x2, p.c1, p.c0);
    const double c2 = GDT_FMA(x2, p.c4, p.c3);
    const double x6 = GDT_MUL(x4, x2);
    const double c = GDT_FMA(x4, p.c2, c1);
    return GDT_FMA(c2, x6, c);
}

GDT_FN uint32_t abstop12(float x) {
    uint32_t u;
#if defined(__CUDA_ARCH__)
    u = __float_as_uint(x);
#else
    memcpy(&u, &x, 4);
#endif
    return (u >> 20) & 0x7ff;
}

// reduce_fast: n = ((int32)(x * (2/pi) * 2^24) + 0x800000) >> 24 ; r = x - n * (pi/2) (fused)
GDT_FN double reduce_fast(double x, int &n) {
    const double r = GDT_MUL(x, 0x1.45f306dc9c883p+23);
    n = ((int32_t)r + 0x800000) >> 24;
    return GDT_FMA(-(double)n, 0x1.921fb54442d18p+0, x);
}

GDT_FN bool in_fast_range(float y) { return abstop12(y) < 0x42f; }   // |y| < 120

GDT_FN float sinf_glibc(float y) {
    const double x = (double)y;
    const uint32_t top = abstop12(y);
    if (top <= 0x3f3) {   // |y| < pi/4 (top12 compare exactly as the library does)
        const double x2 = GDT_MUL(x, x);
        if (top <= 0x397) return y;    // |y| < 2^-12
        return (float)poly_sin(x, x2, tab(0));
    }
    int n;
    const double r = reduce_fast(x, n);
    const Tab p = tab((n & 2) != 0);
    const double r2 = GDT_MUL(r, r);
    if ((n & 1) == 0) {
        const double sign = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
        return (float)poly_sin(GDT_MUL(r, sign), r2, p);
    }
    return (float)poly_cos(r2, p);
}

GDT_FN float cosf_glibc(float y) {
    const double x = (double)y;
    const uint32_t top = abstop12(y);
    if (top <= 0x3f3) {
        const double x2 = GDT_MUL(x, x);
        if (top <= 0x397) return 1.0f;
        return (float)poly_cos(x2, tab(0));
    }
    int n;
    const double r = reduce_fast(x, n);
    const Tab p = tab((n & 2) != 0);
    const double r2 = GDT_MUL(r, r);
    if ((n & 1) != 0) {   // sinf_poly(x * s, x * x, p, n ^ 1)
        const double sign = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
        return (float)poly_sin(GDT_MUL(r, sign), r2, p);
    }
    return (float)poly_cos(r2, p);
}

// sinf and cosf of the same argument in one pass: one range reduction and both polynomials, straight-line
// (selects instead of the library's branches). For |y| < pi/4 the library's short path is the general one with
// n = 0 (r = x exactly, first table, sign +1), so only the |y| < 2^-12 returns need a select. Bit-identical to
// sinf_glibc / cosf_glibc above (tests/csrc/check_trig.cpp sweeps all three against libm).
GDT_FN void sincosf_glibc(float y, float &s_out, float &c_out) {
    const double x = (double)y;
    const uint32_t top = abstop12(y);
    int n;
    const double r = reduce_fast(x, n);
    const Tab p = tab((n & 2) != 0);
    const double r2 = GDT_MUL(r, r);
    const double sign = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    const float ps = (float)poly_sin(GDT_MUL(r, sign), r2, p);
    const float pc = (float)poly_cos(r2, p);
    const bool odd = (n & 1) != 0, tiny = top <= 0x397;
    s_out = tiny ? y : (odd ? pc : ps);
    c_out = tiny ? 1.0f : (odd ? ps : pc);
}

} // namespace gdtrig
