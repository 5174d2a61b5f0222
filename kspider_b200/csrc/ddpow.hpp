// pow(x, e) for 0 < x < 1, 0 < e <= 1 in double-double arithmetic (about 2^-100 relative error before the
// final rounding), so that the result is the correctly rounded double except when the true value lies
// within ~2^-45 ulp of a rounding boundary.  Used by K8 (ANI column) in place of CUDA's pow() (<= 2 ulp):
// CPython's float ** float is glibc's pow (< 1 ulp, correctly rounded in all but a few per cent of a
// per cent of cases), so a correctly rounded value equals the reference's in > 99.9 % of the rows.
// Every operation is an explicit IEEE add / mul / fma (device: __dadd_rn / __dmul_rn / __fma_rn, which the
// compiler never contracts; host build of the same header, for the CPU check of the algorithm:
// -ffp-contract=off).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#ifdef __CUDA_ARCH__
#define KSB_DD_HD __host__ __device__ __forceinline__
#define KSB_DD_ADD(a, b) __dadd_rn((a), (b))
#define KSB_DD_MUL(a, b) __dmul_rn((a), (b))
#define KSB_DD_FMA(a, b, c) __fma_rn((a), (b), (c))
#define KSB_DD_DIV(a, b) __ddiv_rn((a), (b))
#else
#ifdef __CUDACC__
#define KSB_DD_HD __host__ __device__ inline
#else
#define KSB_DD_HD inline
#endif
#define KSB_DD_ADD(a, b) ((a) + (b))
#define KSB_DD_MUL(a, b) ((a) * (b))
#define KSB_DD_FMA(a, b, c) std::fma((a), (b), (c))
#define KSB_DD_DIV(a, b) ((a) / (b))
#endif

namespace ksb {

struct dd { double hi, lo; };

KSB_DD_HD dd dd_two_sum(double a, double b) {
    const double s = KSB_DD_ADD(a, b), bb = KSB_DD_ADD(s, -a);
    const double e = KSB_DD_ADD(KSB_DD_ADD(a, -KSB_DD_ADD(s, -bb)), KSB_DD_ADD(b, -bb));
    return dd{s, e};
}
KSB_DD_HD dd dd_quick_two_sum(double a, double b) {  // |a| >= |b|
    const double s = KSB_DD_ADD(a, b);
    return dd{s, KSB_DD_ADD(b, -KSB_DD_ADD(s, -a))};
}
KSB_DD_HD dd dd_two_prod(double a, double b) {
    const double p = KSB_DD_MUL(a, b);
    return dd{p, KSB_DD_FMA(a, b, -p)};
}
KSB_DD_HD dd dd_add(dd a, dd b) {
    dd s = dd_two_sum(a.hi, b.hi);
    const dd t = dd_two_sum(a.lo, b.lo);
    s.lo = KSB_DD_ADD(s.lo, t.hi);
    s = dd_quick_two_sum(s.hi, s.lo);
    s.lo = KSB_DD_ADD(s.lo, t.lo);
    return dd_quick_two_sum(s.hi, s.lo);
}
KSB_DD_HD dd dd_add_d(dd a, double b) {
    dd s = dd_two_sum(a.hi, b);
    s.lo = KSB_DD_ADD(s.lo, a.lo);
    return dd_quick_two_sum(s.hi, s.lo);
}
KSB_DD_HD dd dd_mul(dd a, dd b) {
    dd p = dd_two_prod(a.hi, b.hi);
    p.lo = KSB_DD_ADD(p.lo, KSB_DD_ADD(KSB_DD_MUL(a.hi, b.lo), KSB_DD_MUL(a.lo, b.hi)));
    return dd_quick_two_sum(p.hi, p.lo);
}
KSB_DD_HD dd dd_mul_d(dd a, double b) {
    dd p = dd_two_prod(a.hi, b);
    p.lo = KSB_DD_ADD(p.lo, KSB_DD_MUL(a.lo, b));
    return dd_quick_two_sum(p.hi, p.lo);
}
KSB_DD_HD dd dd_div(dd a, dd b) {  // three quotient digits
    const double q1 = KSB_DD_DIV(a.hi, b.hi);
    dd r = dd_add(a, dd_mul_d(b, -q1));
    const double q2 = KSB_DD_DIV(r.hi, b.hi);
    r = dd_add(r, dd_mul_d(b, -q2));
    const double q3 = KSB_DD_DIV(r.hi, b.hi);
    dd q = dd_quick_two_sum(q1, q2);
    return dd_add_d(q, q3);
}

// ln(x), x > 0 finite: x = m * 2^n, m in [sqrt(1/2), sqrt(2)); ln m = 2 atanh(t), t = (m - 1) / (m + 1), |t| <= 0.1716
KSB_DD_HD dd dd_log(double x) {
    int n = 0;
    double m = frexp(x, &n);  // [0.5, 1)
    if (m < 0.70710678118654752440) { m = KSB_DD_MUL(m, 2.0); n--; }
    const dd num = dd{KSB_DD_ADD(m, -1.0), 0.0};  // exact (Sterbenz)
    const dd den = dd_two_sum(m, 1.0);
    const dd t = dd_div(num, den);
    const dd t2 = dd_mul(t, t);
    // sum_{k=0}^{17} t^(2k) / (2k + 1), Horner from the top: |t|^2 <= 0.0295, the truncated tail < 2^-100
    const dd inv_odd[18] = {{1.0, 0.0}, {0.3333333333333333, 1.850371707708594e-17}, {0.2, -1.1102230246251566e-17}, {0.14285714285714285, 7.93016446160826e-18}, {0.1111111111111111, 6.1679056923619804e-18}, {0.09090909090909091, -2.523234146875356e-18}, {0.07692307692307693, -4.270088556250602e-18}, {0.06666666666666667, 9.251858538542971e-19}, {0.058823529411764705, 8.163404592832033e-19}, {0.05263157894736842, 2.921639538487254e-18}, {0.047619047619047616, 2.64338815386942e-18}, {0.043478260869565216, 1.206764157201257e-18}, {0.04, -8.326672684688674e-19}, {0.037037037037037035, 2.05596856412066e-18}, {0.034482758620689655, 4.785444071660157e-19}, {0.03225806451612903, 8.953411488912552e-19}, {0.030303030303030304, -8.410780489584519e-19}, {0.02857142857142857, 8.921435019309293e-19}};
    dd acc = inv_odd[17];
#pragma unroll
    for (int k = 16; k >= 0; k--) acc = dd_add(dd_mul(acc, t2), inv_odd[k]);
    dd r = dd_mul(t, acc);
    r = dd_mul_d(r, 2.0);
    const dd ln2 = dd{0.69314718055994530942, 2.3190468138462995584e-17};
    return dd_add(dd_mul_d(ln2, (double)n), r);
}

// exp(z) for -745 < z <= 0 as a double: z = q ln2 + r, |r| <= ln2 / 2; exp(r) = (exp(r / 64))^64 with a 12-term series
KSB_DD_HD double dd_exp_to_double(dd z) {
    const dd ln2 = dd{0.69314718055994530942, 2.3190468138462995584e-17};
    const double qd = nearbyint(KSB_DD_MUL(z.hi, 1.44269504088896340736));
    dd r = dd_add(z, dd_mul_d(ln2, -qd));
    r = dd_mul_d(r, 1.0 / 64.0);  // exact scaling
    // 1 + r + r^2/2! + ... + r^12/12!, |r| <= 0.0055: tail < 2^-120
    const dd inv_fact[13] = {{1.0, 0.0}, {1.0, 0.0}, {0.5, 0.0}, {0.16666666666666666, 9.25185853854297e-18}, {0.041666666666666664, 2.3129646346357427e-18}, {0.008333333333333333, 1.1564823173178714e-19}, {0.001388888888888889, -5.300543954373577e-20}, {0.0001984126984126984, 1.7209558293420705e-22}, {2.48015873015873e-05, 2.1511947866775882e-23}, {2.7557319223985893e-06, -1.858393274046472e-22}, {2.755731922398589e-07, 2.3767714622250297e-23}, {2.505210838544172e-08, -1.448814070935912e-24}, {2.08767569878681e-09, -1.20734505911326e-25}};
    dd acc = inv_fact[12];
#pragma unroll
    for (int k = 11; k >= 0; k--) acc = dd_add(dd_mul(acc, r), inv_fact[k]);
    for (int i = 0; i < 6; i++) acc = dd_mul(acc, acc);
    const double y = KSB_DD_ADD(acc.hi, acc.lo);
    return ldexp(y, (int)qd);
}

// x^e, 0 < x < 1, 0 < e <= 1
KSB_DD_HD double dd_pow(double x, double e) {
    return dd_exp_to_double(dd_mul_d(dd_log(x), e));
}

}  // namespace ksb
