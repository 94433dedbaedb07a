// dependents.cuh -- the per-super-droplet attributes PySDM derives from the water mass for the
// gravitational collision kernels, as device functions shared by the method-level kernels
// (storage_ops.cu, physics.cu), the one-pass kernel pysdm_b200_dependent_attributes and the
// fused collision step (which refreshes them for the few super-droplets a coalescence changed),
// so that every path produces the same bits.
//   volume   = mass / rho_w                       attributes/physics/volume.py:17,
//                                                 impl_numba/methods/physics_methods.py:107-119
//   radius   = (volume * (1 / PI_4_3)) ** (1/3)   attributes/physics/radius.py:15-17
//              (`**=` is sign(x) |x|^e, impl_numba/storage_impl.py `power`)
//   velocity = Gunn-Kinzer table interpolation    dynamics/terminal_velocity/gunn_and_kinzer.py:128-133,
//                                                 impl_numba/methods/terminal_velocity_methods.py:13-30
#pragma once
#include <math.h>

#include "common.cuh"

// Storage.__ipow__: sign(x) * |x| ** e
__device__ __forceinline__ double signed_pow(double x, double e)
{
    double s = x > 0 ? 1.0 : (x < 0 ? -1.0 : (x == 0 ? 0.0 : x));
    double ax = fabs(x);
    double p;
    if (e == 2.0)
        p = ax * ax;  // exact-rounded, what libm's pow returns for squares
    else if (e == 1.0)
        p = ax;
    else
        p = pow(ax, e);
    return s * p;
}

// interpolation body: r_id = int(factor * r); rest = ((factor * r) % 1) / factor;
// out = b[r_id] + rest * c[r_id]; 0 for negative radii
__device__ __forceinline__ double table_interpolation(double r, double factor,
                                                      const double *__restrict__ b,
                                                      const double *__restrict__ c)
{
    if (r < 0) return 0.0;
    double fr = factor * r;
    i64 r_id;
    double frac;  // (factor * r) % 1
    if (fr >= 0.0 && fr < 2251799813685248.0) {  // NaN takes the other branch
        // exact, without the 64-bit conversion and the fmod loop: 2^52 + floor(fr) by a
        // round-down add, the integer from its mantissa, fr - floor(fr) = fmod(fr, 1) exactly
        const double t = __dadd_rd(fr, 4503599627370496.0);
        r_id = (i64)((u64)__double_as_longlong(t) & 0x000FFFFFFFFFFFFFull);
        frac = fr - (t - 4503599627370496.0);
    } else {
        r_id = (i64)fr;
        frac = fmod(fr, 1.0);
    }
    double r_rest = frac / factor;
    return b[r_id] + r_rest * c[r_id];
}

struct DependentParams {
    double rho_w, radius_scale, radius_exponent;  // 1 / PI_4_3, 1 / 3 as the host computed them
    double table_factor, radius_max;              // Interpolation.factor, .maximum_radius
    const double *table_b, *table_c;              // Interpolation.a, .b (601 entries each)
};

// volume, radius, fall velocity of one super-droplet from its water mass; returns false when the
// radius is beyond the table (the reference raises ValueError, gunn_and_kinzer.py:129-133)
__device__ __forceinline__ bool dependents_of(double mass, const DependentParams &p, double &volume,
                                              double &radius, double &velocity)
{
    volume = mass / p.rho_w;
    radius = signed_pow(volume * p.radius_scale, p.radius_exponent);
    const bool ok = !(radius > p.radius_max);
    velocity = ok ? table_interpolation(radius, p.table_factor, p.table_b, p.table_c) : 0.0;
    return ok;
}
