// vt_pow_cr.h — pow(x, y) for x > 0 in double-double arithmetic (~100 bits), rounded once at the end.
//
// lib.material / material_background compute (1 + bounces) ** (1 + falloff) (reference lib.py:450, 465), which is glibc's
// pow in the interpreter.  glibc's pow carries ~68 bits internally and is correctly rounded except when the exact value
// lies within ~0.01 ulp of a rounding boundary; libdevice's pow is within 2 ulp and differed from glibc in the last bit of
// `energy` on 40-80 % of the rays (profiles/r2_parity_counts.json, round-2 first capture).  This function returns the
// correctly rounded value (its own error is ~2^-97, so it misrounds with probability ~2^-44), i.e. glibc's result wherever
// glibc rounds correctly.  The kernel memoises the handful of values a frame asks for (pow_cached), so the ~1 300
// flops here are paid a few hundred times per frame, not per ray.  Host and device compile the same code (the CPU test
// checks it against 60-digit decimal arithmetic, the GPU test checks device == host).
#pragma once
#include <cmath>

#if defined(__CUDA_ARCH__)
#define VTP_FN __device__ __forceinline__
#define VTP_FMA(a, b, c) __fma_rn((a), (b), (c))
#define VTP_MUL(a, b) __dmul_rn((a), (b))
#define VTP_ADD(a, b) __dadd_rn((a), (b))
#elif defined(__CUDACC__)
#define VTP_FN __host__ __device__ inline
#define VTP_FMA(a, b, c) fma((a), (b), (c))
#define VTP_MUL(a, b) ((a) * (b))
#define VTP_ADD(a, b) ((a) + (b))
#else
#define VTP_FN inline
#define VTP_FMA(a, b, c) std::fma((a), (b), (c))
#define VTP_MUL(a, b) ((a) * (b))
#define VTP_ADD(a, b) ((a) + (b))
#endif

namespace vtp {

struct dd { double hi, lo; };

VTP_FN dd two_sum(double a, double b) { double s = VTP_ADD(a, b), bb = VTP_ADD(s, -a); return {s, VTP_ADD(VTP_ADD(a, -VTP_ADD(s, -bb)), VTP_ADD(b, -bb))}; }
VTP_FN dd quick_two_sum(double a, double b) { double s = VTP_ADD(a, b); return {s, VTP_ADD(b, -VTP_ADD(s, -a))}; }
VTP_FN dd two_prod(double a, double b) { double p = VTP_MUL(a, b); return {p, VTP_FMA(a, b, -p)}; }
VTP_FN dd add(dd a, dd b) {
    dd s = two_sum(a.hi, b.hi), t = two_sum(a.lo, b.lo);
    s.lo = VTP_ADD(s.lo, t.hi); s = quick_two_sum(s.hi, s.lo);
    s.lo = VTP_ADD(s.lo, t.lo); return quick_two_sum(s.hi, s.lo);
}
VTP_FN dd add_d(dd a, double b) { dd s = two_sum(a.hi, b); s.lo = VTP_ADD(s.lo, a.lo); return quick_two_sum(s.hi, s.lo); }
VTP_FN dd neg(dd a) { return {-a.hi, -a.lo}; }
VTP_FN dd mul(dd a, dd b) { dd p = two_prod(a.hi, b.hi); p.lo = VTP_ADD(p.lo, VTP_ADD(VTP_MUL(a.hi, b.lo), VTP_MUL(a.lo, b.hi))); return quick_two_sum(p.hi, p.lo); }
VTP_FN dd mul_d(dd a, double b) { dd p = two_prod(a.hi, b); p.lo = VTP_ADD(p.lo, VTP_MUL(a.lo, b)); return quick_two_sum(p.hi, p.lo); }
VTP_FN dd div(dd a, dd b) {
    double q1 = a.hi / b.hi;
    dd r = add(a, neg(mul_d(b, q1)));
    double q2 = r.hi / b.hi;
    r = add(r, neg(mul_d(b, q2)));
    double q3 = r.hi / b.hi;
    return add_d(quick_two_sum(q1, q2), q3);
}

// 1 / (2n + 1) and 1 / n! as double-doubles (generated with exact rational arithmetic).  On the device they live in constant
// memory and the two Horner loops are not unrolled: the function is cold code and must stay small (an unrolled copy with
// the constants as immediates was 1 900 instructions).
#define VTP_ODD_ROWS {0x1.0000000000000p+0, 0x0.0p+0}, {0x1.5555555555555p-2, 0x1.5555555555555p-56}, {0x1.999999999999ap-3, -0x1.999999999999ap-57}, \
        {0x1.2492492492492p-3, 0x1.2492492492492p-57}, {0x1.c71c71c71c71cp-4, 0x1.c71c71c71c71cp-58}, {0x1.745d1745d1746p-4, -0x1.745d1745d1746p-59}, \
        {0x1.3b13b13b13b14p-4, -0x1.3b13b13b13b14p-58}, {0x1.1111111111111p-4, 0x1.1111111111111p-60}, {0x1.e1e1e1e1e1e1ep-5, 0x1.e1e1e1e1e1e1ep-61}, \
        {0x1.af286bca1af28p-5, 0x1.af286bca1af28p-59}, {0x1.8618618618618p-5, 0x1.8618618618618p-59}, {0x1.642c8590b2164p-5, 0x1.642c8590b2164p-60}, \
        {0x1.47ae147ae147bp-5, -0x1.eb851eb851eb8p-61}, {0x1.2f684bda12f68p-5, 0x1.2f684bda12f68p-59}, {0x1.1a7b9611a7b96p-5, 0x1.1a7b9611a7b96p-61}, \
        {0x1.0842108421084p-5, 0x1.0842108421084p-60}, {0x1.f07c1f07c1f08p-6, -0x1.f07c1f07c1f08p-61}, {0x1.d41d41d41d41dp-6, 0x1.0750750750750p-60}, \
        {0x1.bacf914c1bad0p-6, -0x1.bacf914c1bad0p-60}, {0x1.a41a41a41a41ap-6, 0x1.0690690690690p-60}, {0x1.8f9c18f9c18fap-6, -0x1.f3831f3831f38p-61}, \
        {0x1.7d05f417d05f4p-6, 0x1.7d05f417d05f4p-62}, {0x1.6c16c16c16c17p-6, -0x1.f49f49f49f49fp-61}, {0x1.5c9882b931057p-6, 0x1.310572620ae4cp-61}, \
        {0x1.4e5e0a72f0539p-6, 0x1.e0a72f0539783p-60}
#define VTP_FACT_ROWS {0x1.0000000000000p+0, 0x0.0p+0}, {0x1.0000000000000p+0, 0x0.0p+0}, {0x1.0000000000000p-1, 0x0.0p+0}, \
        {0x1.5555555555555p-3, 0x1.5555555555555p-57}, {0x1.5555555555555p-5, 0x1.5555555555555p-59}, {0x1.1111111111111p-7, 0x1.1111111111111p-63}, \
        {0x1.6c16c16c16c17p-10, -0x1.f49f49f49f49fp-65}, {0x1.a01a01a01a01ap-13, 0x1.a01a01a01a01ap-73}, {0x1.a01a01a01a01ap-16, 0x1.a01a01a01a01ap-76}, \
        {0x1.71de3a556c734p-19, -0x1.c154f8ddc6c00p-73}, {0x1.27e4fb7789f5cp-22, 0x1.cbbc05b4fa99ap-76}, {0x1.ae64567f544e4p-26, -0x1.c062e06d1f209p-80}, \
        {0x1.1eed8eff8d898p-29, -0x1.2aec959e14c06p-83}, {0x1.6124613a86d09p-33, 0x1.f28e0cc748ebep-87}, {0x1.93974a8c07c9dp-37, 0x1.05d6f8a2efd1fp-92}, \
        {0x1.ae7f3e733b81fp-41, 0x1.1d8656b0ee8cbp-97}, {0x1.ae7f3e733b81fp-45, 0x1.1d8656b0ee8cbp-101}, {0x1.952c77030ad4ap-49, 0x1.ac981465ddc6cp-103}, \
        {0x1.6827863b97d97p-53, 0x1.eec01221a8b0bp-107}, {0x1.2f49b46814157p-57, 0x1.2650f61dbdcb4p-112}, {0x1.e542ba4020225p-62, 0x1.ea72b4afe3c2fp-120}, \
        {0x1.71b8ef6dcf572p-66, -0x1.d043ae40c4647p-120}, {0x1.0ce396db7f853p-70, -0x1.aebcdbd20331cp-124}, {0x1.761b41316381ap-75, -0x1.3423c7d91404fp-130}, \
        {0x1.f2cf01972f578p-80, -0x1.9ada5fcc1ab14p-135}, {0x1.3f3ccdd165fa9p-84, -0x1.58ddadf344487p-139}, {0x1.88e85fc6a4e5ap-89, -0x1.71c37ebd16540p-143}, \
        {0x1.d1ab1c2dccea3p-94, 0x1.054d0c78aea14p-149}, {0x1.0a18a2635085dp-98, 0x1.b9e2e28e1aa54p-153}, {0x1.259f98b4358adp-103, 0x1.eaf8c39dd9bc5p-157}, \
        {0x1.3932c5047d60ep-108, 0x1.832b7b530a627p-162}
#if defined(__CUDACC__)
__constant__ double ODD_D[25][2] = {VTP_ODD_ROWS};
__constant__ double FACT_D[31][2] = {VTP_FACT_ROWS};
#endif
static const double ODD_H[25][2] = {VTP_ODD_ROWS};
static const double FACT_H[31][2] = {VTP_FACT_ROWS};
#if defined(__CUDA_ARCH__)
#define VTP_ODD ODD_D
#define VTP_FACT FACT_D
#define VTP_NOUNROLL _Pragma("unroll 1")
#else
#define VTP_ODD ODD_H
#define VTP_FACT FACT_H
#define VTP_NOUNROLL
#endif

// pow(x, y), 0 < x < inf, y finite, result in the normal range
VTP_FN double pow_cr(double x, double y) {
    const double LN2_HI = 0x1.62e42fefa39efp-1, LN2_MID = 0x1.abc9e3b39803fp-56, LN2_LO = 0x1.7b57a079a1934p-111;
    const double INV_LN2 = 0x1.71547652b82fep+0;
    // ---- ln x = e ln 2 + 2 atanh((m - 1) / (m + 1)),  x = m 2^e,  m in [sqrt(1/2), sqrt(2))
    int e;
    double m = frexp(x, &e);
    if (m < 0.70710678118654752440) { m *= 2.0; e -= 1; }
    dd s = div({VTP_ADD(m, -1.0), 0.0}, two_sum(m, 1.0));
    dd s2 = mul(s, s);
    dd acc = {VTP_ODD[24][0], VTP_ODD[24][1]};
    VTP_NOUNROLL
    for (int n = 23; n >= 0; n--) acc = add(mul(acc, s2), {VTP_ODD[n][0], VTP_ODD[n][1]});
    dd lnm = mul(s, acc);
    lnm.hi *= 2.0; lnm.lo *= 2.0;
    const double de = (double)e;
    dd lnx = add(add(two_prod(de, LN2_HI), two_prod(de, LN2_MID)), lnm);
    lnx = add_d(lnx, VTP_MUL(de, LN2_LO));
    // ---- exp(y ln x) = 2^k exp(r)
    dd t = mul_d(lnx, y);
    const double k = rint(VTP_MUL(t.hi, INV_LN2));
    dd r = add(t, neg(two_prod(k, LN2_HI)));
    r = add(r, neg(two_prod(k, LN2_MID)));
    r = add_d(r, -VTP_MUL(k, LN2_LO));
    dd ex = {VTP_FACT[30][0], VTP_FACT[30][1]};
    VTP_NOUNROLL
    for (int n = 29; n >= 0; n--) ex = add(mul(ex, r), {VTP_FACT[n][0], VTP_FACT[n][1]});
    return ldexp(VTP_ADD(ex.hi, ex.lo), (int)k);
}

}  // namespace vtp
