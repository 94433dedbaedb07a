// frag.cuh -- device functions of the breakup fragmentation functions shared by the
// method-level kernels (physics.cu) and the fused breakup kernel (breakup_fused.cu).
// Reference: PySDM/backends/impl_numba/methods/fragmentation_methods.py:12-95,218-260,321-378;
// physics/fragmentation_function/straub2010nf.py:13-49; physics/trivia.py:95-108.
#pragma once
#include <math.h>

#include "common.cuh"

constexpr double kPi = 3.14159265358979323846;
constexpr double kCm = 0.01;
constexpr double kSqrt2OverPi = 0.79788456080286535588;  // sqrt(2 / pi)

__device__ __forceinline__ double sq(double x) { return x * x; }
__device__ __forceinline__ double cube(double x) { return x * (x * x); }

// physics/trivia.py:95-108 (Vedder 1987), constants_defaults.py:364-366
__device__ __forceinline__ double erfinv_approx(double c)
{
    const double b = 89.0 / 880.0;
    const double A = 993.0 / 880.0 / 3.0 / b;
    const double A15 = 7.172275788242473;  // pow(A, 1.5) (glibc), a compile-time constant in Numba too
    return 2 * sqrt(A) * sinh(asinh(atanh(c) / 2 / b / A15) / 3);
}

// fragmentation_limiters, fragmentation_methods.py:78-95
struct Limits {
    double vmin, nfmax;
    int has_nfmax;
};
__device__ __forceinline__ void limit(double &n_fragment, double &frag_volume, double x_plus_y,
                                      const Limits &lim)
{
    if (x_plus_y == 0.0) {
        frag_volume = 0.0;
        n_fragment = 1.0;
        return;
    }
    if (isnan(frag_volume) || frag_volume == 0.0) frag_volume = x_plus_y;
    frag_volume = py_min(frag_volume, x_plus_y);
    if (lim.has_nfmax && x_plus_y / frag_volume > lim.nfmax)
        frag_volume = x_plus_y / lim.nfmax;
    else if (frag_volume < lim.vmin)
        frag_volume = x_plus_y;
    n_fragment = x_plus_y / frag_volume;
}

struct StraubResult {
    double Nr1, Nr2, Nr3, Nr4, Nrt, d34;  // in: Nr1..Nr3 as the caller's scratch holds them
    double n_fragment, frag_volume;
};

// Straub 2010 fragment size by inverse-CDF sampling of the four-mode mixture + limiters
// (fragmentation_methods.py:12-50, 218-260, 321-378; straub2010nf.py:13-49)
__device__ __forceinline__ void straub_fragment(double CW, double gam, double ds, double v_max,
                                                double x_plus_y, double u, const Limits &lim,
                                                StraubResult &r)
{
    const double E_D1 = 0.04 * kCm, MU2 = 0.095 * kCm;
    const double E_D1_SQ = 1.6e-07;  // pow(STRAUB_E_D1, 2)
    // straub_Nr (entries that no branch assigns keep the caller's value)
    double Nr1 = r.Nr1, Nr2 = r.Nr2, Nr3 = r.Nr3, Nr4;
    if (gam * CW >= 7.0) Nr1 = fma(gam * CW, 0.088, -0.088 * 7.0);
    if (CW >= 21.0) {
        Nr2 = fma(CW, 0.22, -0.22 * 21.0);
        if (CW <= 46.0) Nr3 = fma(CW, -0.04, 0.04 * 46.0);
    } else {
        Nr3 = 1.0;
    }
    double sigma1 = sqrt(log(CW / 64 / 100 * kCm * kCm / 12 / E_D1_SQ + 1));
    double mu1 = log(E_D1) - sq(sigma1) / 2;
    double sigma2 = py_max(0.0, 7 * (CW - 21) * kCm / 1000) / sqrt(12.0);
    double mu2 = MU2;
    double sigma3 = (1 + 0.76 * sqrt(CW)) * kCm / 100 / sqrt(12.0);
    double mu3 = 0.9 * ds;
    // straub_mass_remainder
    Nr1 = Nr1 * exp(fma(mu1, 3.0, sq(sigma1) * 4.5));
    Nr2 = Nr2 * fma(3 * mu2, sq(sigma2), cube(mu2));
    Nr3 = Nr3 * fma(3 * mu3, sq(sigma3), cube(mu3));
    Nr4 = fma(ds * ds, ds, v_max * (6 / kPi)) - (Nr1 + Nr2 + Nr3);
    double d34;
    if (Nr4 <= 0.0) {
        d34 = 0;
        Nr4 = 0;
    } else {
        d34 = exp(log(Nr4) * (1.0 / 3));
    }
    double Nrt = Nr1 + Nr2 + Nr3 + Nr4;
    double diameter;
    if (Nrt == 0.0) {
        diameter = 0.0;
    } else if (u < Nr1 / Nrt) {
        double X = u * Nrt / Nr1;
        diameter = exp(mu1 + sqrt(2.0) * sigma1 * erfinv_approx(X));
    } else if (u < (Nr2 + Nr1) / Nrt) {
        double X = (u * Nrt - Nr1) / Nr2;
        diameter = mu2 + sqrt(2.0) * sigma2 * erfinv_approx(X);
    } else if (u < (Nr3 + Nr2 + Nr1) / Nrt) {
        double X = (u * Nrt - Nr1 - Nr2) / Nr3;
        diameter = mu3 + sqrt(2.0) * sigma3 * erfinv_approx(X);
    } else {
        diameter = d34;
    }
    double fv = cube(diameter) * kPi / 6, nf;
    limit(nf, fv, x_plus_y, lim);
    r.Nr1 = Nr1;
    r.Nr2 = Nr2;
    r.Nr3 = Nr3;
    r.Nr4 = Nr4;
    r.Nrt = Nrt;
    r.d34 = d34;
    r.n_fragment = nf;
    r.frag_volume = fv;
}
