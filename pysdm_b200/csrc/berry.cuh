// berry.cuh -- Berry (1967) linear collection efficiency of one pair, shared by the
// method-level kernel (physics.cu, pysdm_b200_linear_collection_efficiency) and the fused step
// (fused.cu, kernel 4 "Hydrodynamic").
// Reference: PySDM/backends/impl_numba/methods/collisions_methods.py:743-782; parameters
// (A, B, D1, D2, E1, E2, F1, F2, G1, G2, G3, Mf, Mg): collision_kernels/hydrodynamic.py:11-13.
#pragma once
#include <math.h>

#include "common.cuh"

// Numba's `float ** int` (numba/cpython/numbers.py:214-245): square and multiply
__device__ __forceinline__ double int_power(double a, i64 b)
{
    double r = 1.0;
    bool invert = b < 0;
    i64 e = invert ? -b : b;
    if (e > 0x10000) return pow(a, (double)b);
    while (e != 0) {
        if (e & 1) r *= a;
        e >>= 1;
        a *= a;
    }
    return invert ? 1.0 / r : r;
}

__device__ __forceinline__ double pw(double x, double e, bool is_int)
{
    return is_int ? int_power(x, (i64)e) : pow(x, e);
}

struct Berry {
    double p[13];
    i64 int_mask;  // bit e set: p[e] was passed as an integer (`x ** int` = square and multiply)
};

// efficiency of the pair with radii ra, rb (any order); 0 where the reference leaves its zero fill
__device__ __forceinline__ double berry_efficiency(const Berry &q, double ra, double rb, double unit)
{
    double o = 0.0;
    double r, r_s;
    if (ra > rb) {
        r = ra / unit;
        r_s = rb / unit;
    } else {
        r = rb / unit;
        r_s = ra / unit;
    }
    double p = r_s / r;
    if (p != 0 && p != 1) {
        double G = pw(q.p[8] / r, q.p[12], (q.int_mask >> 12) & 1) + q.p[9] + q.p[10] * r;
        double Gp = pow(1 - p, G);
        if (Gp != 0) {
            double D = q.p[2] / pw(r, q.p[3], (q.int_mask >> 3) & 1);
            double E = q.p[4] / pw(r, q.p[5], (q.int_mask >> 5) & 1);
            double F = pw(q.p[6] / r, q.p[11], (q.int_mask >> 11) & 1) + q.p[7];
            o = py_max(0.0, q.p[0] + q.p[1] * p + D / pow(p, F) + E / Gp);
        }
    }
    return o;
}
