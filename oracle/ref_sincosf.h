/* TEST INFRASTRUCTURE ONLY (oracle/): never linked into the product library.
 *
 * Restatement of glibc 2.39 x86_64 `sinf`/`cosf` (the `__sinf_fma`/`__cosf_fma`
 * ifunc variants that every FMA+AVX2 host selects), which is what the reference's
 * `(float)cos(angle)` / `(float)sin(angle)` resolve to
 * (/root/reference/src/ORBextractor.cc:156-158, built -O3, std::cos(float) -> cosf).
 *
 * Third-party algorithm, not under /root/reference: glibc 2.39
 * sysdeps/ieee754/flt-32/{s_sinf.c,s_cosf.c,s_sincosf.h,s_sincosf_data.c}
 * (ARM optimized-routines sincosf, S. Nagy).  The FMA contraction pattern below
 * was read off the disassembly of /lib/x86_64-linux-gnu/libm.so.6 in this image
 * (every `a + b*c` is a fused vfmadd, the range reduction is one vfnmadd).
 * Pinned: tests/test_sincosf.py compares this against the host libm over EVERY
 * float in [0, 6.5] (1.09e9 values) with zero mismatches.
 * Only the |x| < 120 paths are restated; ORB angles are in [0, 2*pi].
 */
#ifndef ORACLE_REF_SINCOSF_H
#define ORACLE_REF_SINCOSF_H
#include <math.h>
#include <stdint.h>
#include <string.h>

static const double RS_HPI_INV = 0x1.45F306DC9C883p+23; /* 2/pi * 2^24 */
static const double RS_HPI = 0x1.921FB54442D18p0;
static const double RS_C0 = 0x1p0, RS_C1 = -0x1.ffffffd0c621cp-2, RS_C2 = 0x1.55553e1068f19p-5,
                    RS_C3 = -0x1.6c087e89a359dp-10, RS_C4 = 0x1.99343027bf8c3p-16;
static const double RS_S1 = -0x1.555545995a603p-3, RS_S2 = 0x1.1107605230bc4p-7,
                    RS_S3 = -0x1.994eb3774cf24p-13;

static inline uint32_t rs_top12(float f) { uint32_t u; memcpy(&u, &f, 4); return (u >> 20) & 0x7ff; }

/* n even: sine polynomial on (x, x2); n odd: cosine polynomial; neg selects table[1]
 * (cosine coefficients negated, sine coefficients unchanged). */
static inline float rs_poly(double x, double x2, int n, int neg)
{
    if ((n & 1) == 0) {
        double x3 = x * x2;
        double s1 = fma(RS_S3, x2, RS_S2);
        double x7 = x3 * x2;
        double s = fma(x3, RS_S1, x);
        return (float)fma(s1, x7, s);
    } else {
        double sg = neg ? -1.0 : 1.0; /* exact sign flip of every cosine coefficient */
        double x4 = x2 * x2;
        double c2 = fma(sg * RS_C4, x2, sg * RS_C3);
        double c1 = fma(sg * RS_C1, x2, sg * RS_C0);
        double x6 = x4 * x2;
        double c = fma(x4, sg * RS_C2, c1);
        return (float)fma(c2, x6, c);
    }
}

static inline double rs_reduce(double x, int* np)
{
    double r = x * RS_HPI_INV;
    int n = ((int32_t)r + 0x800000) >> 24;
    *np = n;
    return fma(-(double)n, RS_HPI, x);
}

static inline float ref_sinf(float y)
{
    double x = y;
    if (rs_top12(y) < 0x3f4) { /* |y| < pi/4 : abstop12 <= 0x3f3 */
        double x2 = x * x;
        if (rs_top12(y) <= 0x397) return y;
        return rs_poly(x, x2, 0, 0);
    }
    int n;
    static const double sign[4] = {1.0, -1.0, -1.0, 1.0};
    x = rs_reduce(x, &n);
    return rs_poly(x * sign[n & 3], x * x, n, (n & 2) != 0);
}

static inline float ref_cosf(float y)
{
    double x = y;
    if (rs_top12(y) < 0x3f4) {
        double x2 = x * x;
        if (rs_top12(y) <= 0x397) return 1.0f;
        return rs_poly(x, x2, 1, 0);
    }
    int n;
    static const double sign[4] = {1.0, -1.0, -1.0, 1.0};
    x = rs_reduce(x, &n);
    return rs_poly(x * sign[n & 3], x * x, n ^ 1, (n & 2) != 0);
}
#endif
