// physics.cu -- the floating-point rows of the collision path that sit next to the pair
// kernels: terminal velocities (Gunn-Kinzer table
// interpolation, Rogers-Yau, power series), Berry's linear collection efficiency, and the
// breakup fragmentation functions with their limiters.  One launch per reference backend
// method; every kernel is one pass over f64[n] arrays (HBM bound, 8 B per array element).
//
// Reference (paths relative to PySDM/):
//   backends/impl_numba/methods/terminal_velocity_methods.py:13-64
//   backends/impl_numba/methods/collisions_methods.py:743-782
//   backends/impl_numba/methods/fragmentation_methods.py:12-499
//   physics/fragmentation_function/{straub2010nf,lowlist82,feingold1988}.py, physics/trivia.py:95-108
//
// Arithmetic keeps the reference's left-to-right order (-fmad=false).  Where the reference's
// helpers are compiled by Numba with fast-math regardless of Formulae(fastmath=...)
// (conf.JIT_FLAGS: straub_Nr, straub_mass_remainder, ll82_Nr) the explicit fma() forms are
// the ones LLVM produces there; they were identified on the reference's own outputs
// (tests/golden/physics.npz).  Transcendentals come from CUDA's libm (<= 2 ulp).
#include "pysdm_b200.h"
#include <math.h>

#include "common.cuh"
#include "berry.cuh"
#include "dependents.cuh"
#include "frag.cuh"

namespace {

#define TID_OR_RETURN(n)                                        \
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;         \
    if (i >= (n)) return

// ---- terminal velocities ------------------------------------------------------------
__global__ void k_interpolation(double *__restrict__ output, const double *__restrict__ radius,
                                i64 n, double factor, const double *__restrict__ b,
                                const double *__restrict__ c)
{
    TID_OR_RETURN(n);
    output[i] = table_interpolation(radius[i], factor, b, c);
}

__global__ void k_rogers_yau(double *__restrict__ values, const double *__restrict__ radius, i64 n,
                             double small_k, double medium_k, double large_k, double small_r,
                             double medium_r)
{
    TID_OR_RETURN(n);
    double r = radius[i];
    values[i] = r < small_r ? small_k * (r * r)
                            : (r < medium_r ? medium_k * r : large_k * pow(r, 0.5));
}

constexpr int kMaxSeriesTerms = 8;
struct Series {
    double prefactors[kMaxSeriesTerms], powers[kMaxSeriesTerms];
    int n;
};
__global__ void k_power_series(double *__restrict__ values, const double *__restrict__ radius,
                               i64 n, Series s)
{
    TID_OR_RETURN(n);
    double v = 0.0, r = radius[i];
    for (int j = 0; j < s.n; ++j) v = v + s.prefactors[j] * pow(r, s.powers[j] * 3);
    values[i] = v;
}

// ---- linear_collection_efficiency ----------------------------------------------------
__global__ void k_linear_collection_efficiency(Berry q, double *__restrict__ output, i64 out_len,
                                               const double *__restrict__ radii,
                                               const u8 *__restrict__ flag,
                                               const i64 *__restrict__ idx, i64 length, double unit)
{
    // one thread per output slot d = i // 2; at most one of the positions 2d, 2d+1 is a
    // first-in-pair (flags never overlap), the other slots get the reference's zero fill
    TID_OR_RETURN(out_len);
    double o = 0.0;
    for (int s = 0; s < 2; ++s) {
        i64 pos = 2 * i + s;
        if (pos >= length - 1 || !flag[pos]) continue;
        o = berry_efficiency(q, radii[idx[pos]], radii[idx[pos + 1]], unit);
    }
    output[i] = o;
}

// ---- simple fragmentation functions (+ limiters in the same pass) ---------------------
__global__ void k_slams(double *__restrict__ n_fragment, double *__restrict__ frag_volume,
                        const double *__restrict__ x_plus_y, double *__restrict__ probs,
                        const double *__restrict__ rand, i64 n, Limits lim)
{
    TID_OR_RETURN(n);
    double p = 0.0, nf = 1, u = rand[i];
    for (int m = 0; m < 22; ++m) {
        p += 0.91 * pow((double)(m + 2), -1.56);
        if (u < p) {
            nf = m + 2;
            break;
        }
    }
    probs[i] = p;
    double xy = x_plus_y[i], fv = xy / nf;
    limit(nf, fv, xy, lim);
    n_fragment[i] = nf;
    frag_volume[i] = fv;
}

__global__ void k_exp_frag(double *__restrict__ n_fragment, double scale,
                           double *__restrict__ frag_volume, const double *__restrict__ x_plus_y,
                           const double *__restrict__ rand, i64 n, double tol, Limits lim)
{
    TID_OR_RETURN(n);
    double fv = -scale * log(py_max(1 - rand[i], tol)), nf;
    limit(nf, fv, x_plus_y[i], lim);
    n_fragment[i] = nf;
    frag_volume[i] = fv;
}

__global__ void k_feingold_frag(double *__restrict__ n_fragment, double scale,
                                double *__restrict__ frag_volume,
                                const double *__restrict__ x_plus_y,
                                const double *__restrict__ rand, i64 n, double fragtol, Limits lim)
{
    TID_OR_RETURN(n);
    double xy = x_plus_y[i];
    double fv = -scale * log(py_max(1 - rand[i] * scale / xy, fragtol)), nf;
    limit(nf, fv, xy, lim);
    n_fragment[i] = nf;
    frag_volume[i] = fv;
}

__global__ void k_gauss_frag(double *__restrict__ n_fragment, double mu, double sigma,
                             double *__restrict__ frag_volume, const double *__restrict__ x_plus_y,
                             const double *__restrict__ rand, i64 n, Limits lim)
{
    TID_OR_RETURN(n);
    double fv = mu + sigma * erfinv_approx(rand[i]), nf;
    limit(nf, fv, x_plus_y[i], lim);
    n_fragment[i] = nf;
    frag_volume[i] = fv;
}

__global__ void k_ll82_coalescence_check(double *__restrict__ Ec, const double *__restrict__ dl,
                                         i64 n)
{
    TID_OR_RETURN(n);
    if (dl[i] < 0.4e-3) Ec[i] = 1.0;
}

// ---- Straub 2010 (fragmentation_methods.py:12-50, 321-378; straub2010nf.py:13-49) --------
__global__ void k_straub_frag(double *__restrict__ n_fragment, const double *__restrict__ CW_,
                              const double *__restrict__ gam_, const double *__restrict__ ds_,
                              double *__restrict__ frag_volume, const double *__restrict__ v_max,
                              const double *__restrict__ x_plus_y, const double *__restrict__ rand,
                              double *__restrict__ Nr1_, double *__restrict__ Nr2_,
                              double *__restrict__ Nr3_, double *__restrict__ Nr4_,
                              double *__restrict__ Nrt_, double *__restrict__ d34_, i64 n,
                              Limits lim)
{
    TID_OR_RETURN(n);
    // straub_Nr: the scratch arrays are read-modify-write in the reference (entries that no
    // branch assigns keep the caller's value), so load them first
    StraubResult r;
    r.Nr1 = Nr1_[i];
    r.Nr2 = Nr2_[i];
    r.Nr3 = Nr3_[i];
    straub_fragment(CW_[i], gam_[i], ds_[i], v_max[i], x_plus_y[i], rand[i], lim, r);
    const double Nr1 = r.Nr1, Nr2 = r.Nr2, Nr3 = r.Nr3, Nr4 = r.Nr4, Nrt = r.Nrt, d34 = r.d34;
    const double nf = r.n_fragment, fv = r.frag_volume;
    Nr1_[i] = Nr1;
    Nr2_[i] = Nr2;
    Nr3_[i] = Nr3;
    Nr4_[i] = Nr4;
    Nrt_[i] = Nrt;
    d34_[i] = d34;
    n_fragment[i] = nf;
    frag_volume[i] = fv;
}

// ---- Low & List 1982 (lowlist82.py; fragmentation_methods.py:52-74, 380-472) -------------
struct Hms {
    double H, mu, sigma;
};

__device__ __forceinline__ double ll82_sigma_iter(double H, double d_minus_mu)
{
    double sigma = 1 / H;
    for (int it = 0; it < 10; ++it)
        sigma = 1 / H * kSqrt2OverPi / (1 + erf(d_minus_mu / (sqrt(2.0) * sigma)));
    return sigma;
}

__device__ Hms ll82_f3(double dsCM, double dlCM)  // lowlist82.py:40-98
{
    double Ff1 = py_max(0.0, (-2.25e4 * sq(dlCM - 0.403) - 37.9) * pow(dsCM, 2.5) +
                                 9.67 * sq(dlCM - 0.170) + 4.95);
    double Ff2 = 1.02e4 * pow(dsCM, 2.83) + 2;
    double ds0 = py_max(0.04, pow(Ff1 / 2.83, 1 / 1.02e4));
    double Ff = dsCM > ds0 ? py_max(2.0, Ff1) : py_max(2.0, Ff2);
    double Dff3 = 0.241 * dsCM + 0.0129;
    double Pf301 = 1.68e5 * pow(dsCM, 2.33);
    double Pf302 =
        py_max(0.0, (43.4 * sq(dlCM + 1.81) - 159.0) / dsCM - 3870 * sq(dlCM - 0.285) - 58.1);
    double alpha = (dsCM - ds0) / (0.2 * ds0);
    double Pf303 = alpha * Pf301 + (1 - alpha) * Pf302;
    double Pf0 = dsCM < ds0 ? Pf301 : (dsCM > 1.2 * ds0 ? Pf302 : Pf303);
    Hms r;
    r.sigma = 10 * Dff3;
    r.mu = log(Dff3) + sq(r.sigma);
    r.H = Pf0 * Dff3 / exp(-0.5 * sq(r.sigma));
    for (int it = 0; it < 10; ++it) {
        if (r.sigma == 0.0 || r.H == 0) return Hms{0.0, log(ds0), log(ds0)};
        r.sigma = kSqrt2OverPi * (Ff - 2) / r.H /
                  (1 - erf((log(0.01) - r.mu) / sqrt(2.0) / r.sigma));
        r.mu = log(Dff3) + sq(r.sigma);
        r.H = Pf0 * Dff3 / exp(-0.5 * sq(r.sigma));
    }
    return r;
}

__device__ Hms ll82_s2(double dlCM, double dsCM, double St)  // :119-142
{
    double Dss2 = 0.254 * pow(dsCM, 0.413) * exp(3.53 * pow(dsCM, 2.51) * (dlCM - dsCM));
    double bstar = 14.2 * exp(-17.2 * dsCM);
    double Ps20 = 0.23 * pow(dsCM, -3.93) * pow(dlCM, bstar);
    Hms r;
    r.sigma = 10 * Dss2;
    r.mu = log(Dss2) + sq(r.sigma);
    r.H = Ps20 * Dss2 / exp(-0.5 * sq(r.sigma));
    double Fs = 5 * erf((St - 2.52e-6) / (1.85e-6)) + 6;
    for (int it = 0; it < 10; ++it) {
        r.sigma = kSqrt2OverPi * (Fs - 1) / r.H /
                  (1 - erf((log(0.01) - r.mu) / sqrt(2.0) / r.sigma));
        r.mu = log(Dss2) + sq(r.sigma);
        r.H = Ps20 * Dss2 / exp(-0.5 * sq(r.sigma));
    }
    return r;
}

__device__ Hms ll82_d2(double dsCM, double dlCM, double CKE)  // :160-190
{
    double Ddd2 = exp(-17.4 * dsCM - 0.671 * (dlCM - dsCM)) * dsCM;
    double bstar = 0.007 * pow(dsCM, -2.54);
    double Pd20 = 0.0884 * pow(dsCM, -2.52) * pow(dlCM - dsCM, bstar);
    Hms none{0.0, log(Ddd2), log(Ddd2)};
    Hms r;
    r.sigma = 10 * Ddd2;
    r.mu = log(Ddd2) + sq(r.sigma);
    r.H = Pd20 * Ddd2 / exp(-0.5 * sq(r.sigma));
    double Fd = py_max(1.0, 297.5 + 23.7 * log(CKE));
    if (Fd == 1.0) return none;
    for (int it = 0; it < 10; ++it) {
        if (r.sigma == 0.0 || r.H <= 0.1) return none;
        if (r.sigma >= 1.0) return none;
        r.sigma = kSqrt2OverPi * (Fd - 1) / r.H /
                  (1 - erf((log(0.01) - r.mu) / sqrt(2.0) / r.sigma));
        r.mu = log(Ddd2) + sq(r.sigma);
        r.H = Pd20 * Ddd2 / exp(-0.5 * sq(r.sigma));
    }
    return r;
}

__device__ __forceinline__ double ll82_normal(const Hms &m, double X)
{
    return m.mu + sqrt(2.0) * m.sigma * erfinv_approx(2 * X - 1);
}

__global__ void k_ll82_frag(double *__restrict__ n_fragment, const double *__restrict__ CKE_,
                            const double *__restrict__ W_, const double *__restrict__ W2_,
                            const double *__restrict__ St_, const double *__restrict__ ds_,
                            const double *__restrict__ dl_, const double *__restrict__ dcoal_,
                            double *__restrict__ frag_volume, const double *__restrict__ x_plus_y,
                            double *__restrict__ rand, double *__restrict__ Rf_,
                            double *__restrict__ Rs_, double *__restrict__ Rd_, i64 n, double tol,
                            Limits lim)
{
    TID_OR_RETURN(n);
    const double dl = dl_[i], ds = ds_[i], dcoal = dcoal_[i];
    double fv, nf;
    if (dl <= 0.4e-3) {
        fv = cube(dcoal) * kPi / 6;
    } else if (ds == 0.0 || dl == 0.0) {
        fv = 1e-18;
    } else {
        const double CKE = CKE_[i], W = W_[i];
        const double dsCM = ds / kCm, dlCM = dl / kCm, dcoalCM = dcoal / kCm;
        // ll82_Nr (fast-math helper in the reference: see the file header)
        double Rf = CKE >= 0.893e-6 ? 1.11e-4 * pow(CKE, -0.654) : 1.0;
        double Rs = W >= 0.86 ? fma(-0.685, exp(fma(W2_[i], -1.63, 1.63 * 0.86)), 0.685) : 0.0;
        double Rd = (Rs + Rf) > 1.0 ? 0.0 : 1.0 - (Rs + Rf);
        Rf_[i] = Rf;
        Rs_[i] = Rs;
        Rd_[i] = Rd;
        double u = rand[i];
        if (u <= Rf) {  // filament breakup
            Hms m1, m2, m3 = ll82_f3(dsCM, dlCM);
            m1.H = 50.8 * pow(dlCM, -0.718);  // params_f1, lowlist82.py:15-30
            m1.mu = dlCM;
            m1.sigma = ll82_sigma_iter(m1.H, dcoalCM - dlCM);
            m2.H = 4.18 * pow(dsCM, -1.17);  // params_f2, :32-38
            m2.mu = dsCM;
            m2.sigma = 1 / (sqrt(2 * kPi) * m2.H);
            double H1 = m1.H * m1.mu, H2 = m2.H * m2.mu, H3 = m3.H * exp(m3.mu);
            double Hsum = H1 + H2 + H3;
            u = u / Rf;
            if (u <= H1 / Hsum)
                fv = ll82_normal(m1, py_max(u * Hsum / H1, tol));
            else if (u <= (H1 + H2) / Hsum)
                fv = ll82_normal(m2, (u * Hsum - H1) / H2);
            else
                fv = exp(ll82_normal(m3, py_min((u * Hsum - H1 - H2) / H3, 1.0 - tol)));
        } else if (u <= Rf + Rs) {  // sheet breakup
            Hms m1, m2 = ll82_s2(dlCM, dsCM, St_[i]);
            m1.H = 100 * exp(-3.25 * dsCM);  // params_s1, :100-117
            m1.mu = dlCM;
            m1.sigma = ll82_sigma_iter(m1.H, dcoalCM - dlCM);
            double H1 = m1.H * m1.mu, H2 = m2.H * exp(m2.mu);
            double Hsum = H1 + H2;
            u = (u - Rf) / Rs;
            if (u <= H1 / Hsum)
                fv = ll82_normal(m1, py_max(u * Hsum / H1, tol));
            else
                fv = exp(ll82_normal(m2, py_min((u * Hsum - H1) / H2, 1.0 - tol)));
        } else {  // disk breakup
            Hms m1, m2 = ll82_d2(dsCM, dlCM, CKE);
            m1.mu = dlCM * (1 - exp(-3.70 * (3.10 - W)));  // params_d1, :144-158
            m1.H = 1.58e-5 * pow(CKE, -1.22);
            m1.sigma = ll82_sigma_iter(m1.H, dcoalCM - m1.mu);
            double H1 = m1.H * m1.mu, H2 = m2.H;
            double Hsum = H1 + H2;
            u = (u - Rf - Rs) / Rd;
            if (u <= H1 / Hsum)
                fv = ll82_normal(m1, py_max(u * Hsum / H1, tol));
            else
                fv = exp(ll82_normal(m2, py_min((u * Hsum - H1) / H2, 1 - tol)));
        }
        rand[i] = u;  // the reference rescales rand in place
        fv = fv * 0.01;  // diameter in cm -> m
        fv = cube(fv) * kPi / 6;
    }
    limit(nf, fv, x_plus_y[i], lim);
    n_fragment[i] = nf;
    frag_volume[i] = fv;
}

}  // namespace

// ---- C-ABI -----------------------------------------------------------------------------
#define ST ((cudaStream_t)stream)
#define GRID(n) grid_for(n), kThreads, 0, ST

PYSDM_EXPORT int pysdm_b200_interpolation(double *output, const double *radius, i64 n,
                                          double factor, const double *b, const double *c,
                                          void *stream)
{
    if (n <= 0) return 0;
    k_interpolation<<<GRID(n)>>>(output, radius, n, factor, b, c);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_terminal_velocity_rogers_yau(double *values, const double *radius,
                                                         i64 n, double small_k, double medium_k,
                                                         double large_k, double small_r_limit,
                                                         double medium_r_limit, void *stream)
{
    if (n <= 0) return 0;
    k_rogers_yau<<<GRID(n)>>>(values, radius, n, small_k, medium_k, large_k, small_r_limit,
                              medium_r_limit);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_power_series(double *values, const double *radius, i64 n,
                                         i64 num_terms, const double *host_prefactors,
                                         const double *host_powers, void *stream)
{
    if (n <= 0) return 0;
    if (num_terms < 0 || num_terms > kMaxSeriesTerms) return (int)cudaErrorInvalidValue;
    Series s;
    s.n = (int)num_terms;
    for (int j = 0; j < s.n; ++j) {
        s.prefactors[j] = host_prefactors[j];
        s.powers[j] = host_powers[j];
    }
    k_power_series<<<GRID(n)>>>(values, radius, n, s);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_linear_collection_efficiency(const double *host_params13,
                                                         i64 int_mask, double *output,
                                                         i64 out_len, const double *radii,
                                                         const u8 *is_first_in_pair,
                                                         const i64 *idx, i64 length, double unit,
                                                         void *stream)
{
    if (out_len <= 0) return 0;
    Berry q;
    for (int e = 0; e < 13; ++e) q.p[e] = host_params13[e];
    q.int_mask = int_mask;
    k_linear_collection_efficiency<<<GRID(out_len)>>>(q, output, out_len, radii, is_first_in_pair,
                                                      idx, length, unit);
    return LAUNCH_CHECK();
}

static inline Limits limits_of(double vmin, int has_nfmax, double nfmax)
{
    Limits l;
    l.vmin = vmin;
    l.nfmax = nfmax;
    l.has_nfmax = has_nfmax;
    return l;
}

PYSDM_EXPORT int pysdm_b200_slams_fragmentation(double *n_fragment, double *frag_volume,
                                                const double *x_plus_y, double *probs,
                                                const double *rand, i64 n, double vmin,
                                                int has_nfmax, double nfmax, void *stream)
{
    if (n <= 0) return 0;
    k_slams<<<GRID(n)>>>(n_fragment, frag_volume, x_plus_y, probs, rand, n,
                         limits_of(vmin, has_nfmax, nfmax));
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_exp_fragmentation(double *n_fragment, double scale,
                                              double *frag_volume, const double *x_plus_y,
                                              const double *rand, i64 n, double vmin,
                                              int has_nfmax, double nfmax, double tol,
                                              void *stream)
{
    if (n <= 0) return 0;
    k_exp_frag<<<GRID(n)>>>(n_fragment, scale, frag_volume, x_plus_y, rand, n, tol,
                            limits_of(vmin, has_nfmax, nfmax));
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_feingold1988_fragmentation(double *n_fragment, double scale,
                                                       double *frag_volume,
                                                       const double *x_plus_y, const double *rand,
                                                       i64 n, double fragtol, double vmin,
                                                       int has_nfmax, double nfmax, void *stream)
{
    if (n <= 0) return 0;
    k_feingold_frag<<<GRID(n)>>>(n_fragment, scale, frag_volume, x_plus_y, rand, n, fragtol,
                                 limits_of(vmin, has_nfmax, nfmax));
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_gauss_fragmentation(double *n_fragment, double mu, double sigma,
                                                double *frag_volume, const double *x_plus_y,
                                                const double *rand, i64 n, double vmin,
                                                int has_nfmax, double nfmax, void *stream)
{
    if (n <= 0) return 0;
    k_gauss_frag<<<GRID(n)>>>(n_fragment, mu, sigma, frag_volume, x_plus_y, rand, n,
                              limits_of(vmin, has_nfmax, nfmax));
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_straub_fragmentation(
    double *n_fragment, const double *CW, const double *gam, const double *ds, double *frag_volume,
    const double *v_max, const double *x_plus_y, const double *rand, double *Nr1, double *Nr2,
    double *Nr3, double *Nr4, double *Nrt, double *d34, i64 n, double vmin, int has_nfmax,
    double nfmax, void *stream)
{
    if (n <= 0) return 0;
    k_straub_frag<<<GRID(n)>>>(n_fragment, CW, gam, ds, frag_volume, v_max, x_plus_y, rand, Nr1,
                               Nr2, Nr3, Nr4, Nrt, d34, n, limits_of(vmin, has_nfmax, nfmax));
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_ll82_fragmentation(
    double *n_fragment, const double *CKE, const double *W, const double *W2, const double *St,
    const double *ds, const double *dl, const double *dcoal, double *frag_volume,
    const double *x_plus_y, double *rand, double *Rf, double *Rs, double *Rd, i64 n, double tol,
    double vmin, int has_nfmax, double nfmax, void *stream)
{
    if (n <= 0) return 0;
    k_ll82_frag<<<GRID(n)>>>(n_fragment, CKE, W, W2, St, ds, dl, dcoal, frag_volume, x_plus_y,
                             rand, Rf, Rs, Rd, n, tol, limits_of(vmin, has_nfmax, nfmax));
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_ll82_coalescence_check(double *Ec, const double *dl, i64 n,
                                                   void *stream)
{
    if (n <= 0) return 0;
    k_ll82_coalescence_check<<<GRID(n)>>>(Ec, dl, n);
    return LAUNCH_CHECK();
}
