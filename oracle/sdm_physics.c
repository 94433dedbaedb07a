/*
 * sdm_physics.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see sdm_oracle.c).
 *
 * Plain-C restatement of the floating-point "physics" rows of the collision path
 * (SURVEY.md 8a #11-#14, #25, #26): Storage `**=`, Gunn-Kinzer interpolation,
 * Rogers-Yau terminal velocity, power series, Berry's linear collection efficiency,
 * the fragmentation functions with their limiters and ll82_coalescence_check.
 * Each function cites the reference file:line it follows (paths relative to
 * /root/reference/PySDM/).  Pinned against tests/golden/physics.npz, which
 * oracle/gen_golden_physics.py produced by running the reference's Numba backend
 * (Formulae(fastmath=False)) in the build container.
 *
 * Expressions keep the reference's left-to-right evaluation order; integer powers
 * follow Numba's lowering (numba/cpython/numbers.py:206-310: square-and-multiply
 * for `float ** int`), real powers/transcendentals call libm as Numba does.
 */
#include <math.h>
#include <stdint.h>

typedef int64_t i64;
typedef uint8_t u8;
#define ORACLE_API __attribute__((visibility("default")))

static const double PI = 3.14159265358979323846;
static const double CM = 0.01;

static inline double sq(double x) { return x * x; }
static inline double cube(double x) { return x * (x * x); } /* static_power_impl: res=x; val=x*x; res*=val */
static inline double dmax(double a, double b) { return b > a ? b : a; } /* Python max(a, b) */
static inline double dmin(double a, double b) { return b < a ? b : a; } /* Python min(a, b) */

/* numba int_power_impl (numbers.py:214-245): float ** runtime int */
static inline double int_power(double a, i64 b)
{
    double r = 1.0;
    int invert = b < 0;
    i64 e = invert ? -b : b;
    if (e > 0x10000) return pow(a, (double)b);
    while (e != 0) {
        if (e & 1) r *= a;
        e >>= 1;
        a *= a;
    }
    return invert ? 1.0 / r : r;
}

/* ---- Storage.__ipow__: backends/impl_numba/storage_impl.py:75-78 -------------------
 * output[:] = np.sign(output) * np.power(np.abs(output), exponent) */
ORACLE_API void oracle_power(double *x, i64 n, double e)
{
    for (i64 i = 0; i < n; ++i) {
        double v = x[i];
        double s = v > 0 ? 1.0 : (v < 0 ? -1.0 : (v == 0 ? 0.0 : v));
        x[i] = s * pow(fabs(v), e);
    }
}

/* ---- Storage.product (storage_impl.py multiply_out_of_place), used by
 * attributes/physics/radius.py:15-17 */
ORACLE_API void oracle_product_scalar(double *out, const double *in, double s, i64 n)
{
    for (i64 i = 0; i < n; ++i) out[i] = in[i] * s;
}

/* ---- interpolation: backends/impl_numba/methods/terminal_velocity_methods.py:13-30 */
ORACLE_API void oracle_interpolation(double *output, const double *radius, i64 n, double factor,
                                     const double *b, const double *c)
{
    for (i64 i = 0; i < n; ++i) {
        if (radius[i] < 0) {
            output[i] = 0;
        } else {
            double fr = factor * radius[i];
            i64 r_id = (i64)fr;
            double r_rest = fmod(fr, 1.0) / factor; /* fr >= 0: Python % == fmod */
            output[i] = b[r_id] + r_rest * c[r_id];
        }
    }
}

/* ---- Rogers-Yau: physics/terminal_velocity/rogers_yau.py:13-23 via
 * terminal_velocity_methods.py:32-44.  params = (small_k, medium_k, large_k,
 * small_r_limit, medium_r_limit) */
ORACLE_API void oracle_terminal_velocity_rogers_yau(double *values, const double *radius, i64 n,
                                                    const double *params)
{
    for (i64 i = 0; i < n; ++i) {
        double r = radius[i];
        double small = params[0] * (r * r);
        double medium = params[1] * r;
        double large = params[2] * pow(r, 0.5);
        values[i] = r < params[3] ? small : (r < params[4] ? medium : large);
    }
}

/* ---- power_series: terminal_velocity_methods.py:46-64 */
ORACLE_API void oracle_power_series(double *values, const double *radius, i64 n, i64 num_terms,
                                    const double *prefactors, const double *powers)
{
    for (i64 i = 0; i < n; ++i) {
        values[i] = 0.0;
        for (i64 j = 0; j < num_terms; ++j)
            values[i] = values[i] + prefactors[j] * pow(radius[i], powers[j] * 3);
    }
}

/* ---- linear_collection_efficiency: backends/impl_numba/methods/collisions_methods.py:743-782.
 * int_mask bit e set = params[e] is a Python int (Numba then lowers `x ** params[e]`
 * as square-and-multiply instead of pow): relevant for D2 (3), E2 (5), Mf (11), Mg (12) */
static inline double pw(double x, double e, int is_int)
{
    return is_int ? int_power(x, (i64)e) : pow(x, e);
}

ORACLE_API void oracle_linear_collection_efficiency(const double *params, i64 int_mask,
                                                    double *output, i64 out_len,
                                                    const double *radii, const u8 *flag,
                                                    const i64 *idx, i64 length, double unit)
{
    const double A = params[0], B = params[1], D1 = params[2], D2 = params[3], E1 = params[4],
                 E2 = params[5], F1 = params[6], F2 = params[7], G1 = params[8], G2 = params[9],
                 G3 = params[10], Mf = params[11], Mg = params[12];
    for (i64 i = 0; i < out_len; ++i) output[i] = 0;
    for (i64 i = 0; i < length - 1; ++i) {
        if (!flag[i]) continue;
        double r, r_s;
        if (radii[idx[i]] > radii[idx[i + 1]]) {
            r = radii[idx[i]] / unit;
            r_s = radii[idx[i + 1]] / unit;
        } else {
            r = radii[idx[i + 1]] / unit;
            r_s = radii[idx[i]] / unit;
        }
        double p = r_s / r;
        if (p != 0 && p != 1) {
            double G = pw(G1 / r, Mg, (int)(int_mask >> 12) & 1) + G2 + G3 * r;
            double Gp = pow(1 - p, G);
            if (Gp != 0) {
                double D = D1 / pw(r, D2, (int)(int_mask >> 3) & 1);
                double E = E1 / pw(r, E2, (int)(int_mask >> 5) & 1);
                double F = pw(F1 / r, Mf, (int)(int_mask >> 11) & 1) + F2;
                double o = A + B * p + D / pow(p, F) + E / Gp;
                output[i / 2] = dmax(0, o);
            }
        }
    }
}

/* ---- physics/trivia.py:95-108 erfinv_approx (Vedder 1987), constants
 * physics/constants_defaults.py:364-366 */
static inline double erfinv_approx(double c)
{
    const double b = 89.0 / 880.0;
    const double A = 993.0 / 880.0 / 3.0 / b;
    return 2 * sqrt(A) * sinh(asinh(atanh(c) / 2 / b / pow(A, 3.0 / 2.0)) / 3);
}

/* ---- fragmentation_limiters: backends/impl_numba/methods/fragmentation_methods.py:78-95.
 * has_nfmax = 0 encodes `nfmax is None` */
ORACLE_API void oracle_fragmentation_limiters(double *n_fragment, double *frag_volume, i64 n,
                                              double vmin, int has_nfmax, double nfmax,
                                              const double *x_plus_y)
{
    for (i64 i = 0; i < n; ++i) {
        if (x_plus_y[i] == 0.0) {
            frag_volume[i] = 0.0;
            n_fragment[i] = 1.0;
        } else {
            if (isnan(frag_volume[i]) || frag_volume[i] == 0.0) frag_volume[i] = x_plus_y[i];
            frag_volume[i] = dmin(frag_volume[i], x_plus_y[i]);
            if (has_nfmax && x_plus_y[i] / frag_volume[i] > nfmax)
                frag_volume[i] = x_plus_y[i] / nfmax;
            else if (frag_volume[i] < vmin)
                frag_volume[i] = x_plus_y[i];
            n_fragment[i] = x_plus_y[i] / frag_volume[i];
        }
    }
}

/* ---- slams: fragmentation_methods.py:106-119 (limiters applied by the caller) */
ORACLE_API void oracle_slams_fragmentation(double *n_fragment, double *frag_volume,
                                           const double *x_plus_y, double *probs,
                                           const double *rand, i64 n)
{
    for (i64 i = 0; i < n; ++i) {
        probs[i] = 0.0;
        n_fragment[i] = 1;
        for (int m = 0; m < 22; ++m) {
            probs[i] += 0.91 * pow((double)(m + 2), -1.56);
            if (rand[i] < probs[i]) {
                n_fragment[i] = m + 2;
                break;
            }
        }
        frag_volume[i] = x_plus_y[i] / n_fragment[i];
    }
}

/* ---- exponential: fragmentation_methods.py:136-143 */
ORACLE_API void oracle_exp_fragmentation(double scale, double *frag_volume, const double *rand,
                                         i64 n, double tol)
{
    for (i64 i = 0; i < n; ++i) frag_volume[i] = -scale * log(dmax(1 - rand[i], tol));
}

/* ---- Feingold 1988: fragmentation_methods.py:486-497,
 * physics/fragmentation_function/feingold1988.py:12-14 */
ORACLE_API void oracle_feingold1988_fragmentation(double scale, double *frag_volume,
                                                  const double *x_plus_y, const double *rand,
                                                  i64 n, double fragtol)
{
    for (i64 i = 0; i < n; ++i)
        frag_volume[i] = -scale * log(dmax(1 - rand[i] * scale / x_plus_y[i], fragtol));
}

/* ---- Gaussian: fragmentation_methods.py:474-483 */
ORACLE_API void oracle_gauss_fragmentation(double mu, double sigma, double *frag_volume,
                                           const double *rand, i64 n)
{
    for (i64 i = 0; i < n; ++i) frag_volume[i] = mu + sigma * erfinv_approx(rand[i]);
}

/* ---- ll82_coalescence_check: fragmentation_methods.py:305-319 */
ORACLE_API void oracle_ll82_coalescence_check(double *Ec, const double *dl, i64 n)
{
    for (i64 i = 0; i < n; ++i)
        if (dl[i] < 0.4e-3) Ec[i] = 1.0;
}

/* ---- Straub 2010: fragmentation_methods.py:12-50 (straub_Nr, straub_mass_remainder),
 * :321-378 (body); parameters physics/fragmentation_function/straub2010nf.py:13-49,
 * constants physics/constants_defaults.py:359-361 */
ORACLE_API void oracle_straub_fragmentation(const double *CW, const double *gam, const double *ds,
                                            const double *v_max, double *frag_volume,
                                            const double *rand, double *Nr1, double *Nr2,
                                            double *Nr3, double *Nr4, double *Nrt, double *d34,
                                            i64 n)
{
    const double E_D1 = 0.04 * CM, MU2 = 0.095 * CM;
    for (i64 i = 0; i < n; ++i) {
        /* straub_Nr.  The two helpers are jitted with conf.JIT_FLAGS (fastmath=True whatever
         * Formulae says, impl_numba/conf.py:5-10): LLVM distributes the constant factors and
         * contracts into FMAs.  The forms below are the ones that reproduce the reference's
         * output bits (found by replaying tests/golden/physics.npz). */
        if (gam[i] * CW[i] >= 7.0) Nr1[i] = fma(gam[i] * CW[i], 0.088, -0.088 * 7.0);
        if (CW[i] >= 21.0) {
            Nr2[i] = fma(CW[i], 0.22, -0.22 * 21.0);
            if (CW[i] <= 46.0) Nr3[i] = fma(CW[i], -0.04, 0.04 * 46.0);
        } else {
            Nr3[i] = 1.0;
        }
        Nr4[i] = 1.0;
        Nrt[i] = Nr1[i] + Nr2[i] + Nr3[i] + Nr4[i];
        /* parameters */
        double sigma1 = sqrt(log(CW[i] / 64 / 100 * CM * CM / 12 / pow(E_D1, 2.0) + 1));
        double mu1 = log(E_D1) - sq(sigma1) / 2;
        double sigma2 = dmax(0.0, 7 * (CW[i] - 21) * CM / 1000) / sqrt(12.0);
        double mu2 = MU2;
        double sigma3 = (1 + 0.76 * sqrt(CW[i])) * CM / 100 / sqrt(12.0);
        double mu3 = 0.9 * ds[i];
        /* straub_mass_remainder (fast-math helper, see above) */
        Nr1[i] = Nr1[i] * exp(fma(mu1, 3.0, sq(sigma1) * 4.5));
        Nr2[i] = Nr2[i] * fma(3 * mu2, sq(sigma2), cube(mu2));
        Nr3[i] = Nr3[i] * fma(3 * mu3, sq(sigma3), cube(mu3));
        Nr4[i] = fma(ds[i] * ds[i], ds[i], v_max[i] * (6 / PI)) - (Nr1[i] + Nr2[i] + Nr3[i]);
        if (Nr4[i] <= 0.0) {
            d34[i] = 0;
            Nr4[i] = 0;
        } else {
            d34[i] = exp(log(Nr4[i]) * (1.0 / 3));
        }
        Nrt[i] = Nr1[i] + Nr2[i] + Nr3[i] + Nr4[i];
        double diameter;
        if (Nrt[i] == 0.0) {
            diameter = 0.0;
        } else if (rand[i] < Nr1[i] / Nrt[i]) {
            double X = rand[i] * Nrt[i] / Nr1[i];
            diameter = exp(mu1 + sqrt(2.0) * sigma1 * erfinv_approx(X));
        } else if (rand[i] < (Nr2[i] + Nr1[i]) / Nrt[i]) {
            double X = (rand[i] * Nrt[i] - Nr1[i]) / Nr2[i];
            diameter = mu2 + sqrt(2.0) * sigma2 * erfinv_approx(X);
        } else if (rand[i] < (Nr3[i] + Nr2[i] + Nr1[i]) / Nrt[i]) {
            double X = (rand[i] * Nrt[i] - Nr1[i] - Nr2[i]) / Nr3[i];
            diameter = mu3 + sqrt(2.0) * sigma3 * erfinv_approx(X);
        } else {
            diameter = d34[i];
        }
        frag_volume[i] = cube(diameter) * PI / 6;
    }
}

/* ---- Low & List 1982: physics/fragmentation_function/lowlist82.py (params_*),
 * fragmentation_methods.py:52-74 (ll82_Nr), :380-472 (body) */
typedef struct { double H, mu, sigma; } hms;

static const double SQRT_2_OVER_PI = 0.79788456080286535588; /* np.sqrt(2 / np.pi) */

/* lowlist82.py:15-30 (f1), :101-117 (s1), :144-158 (d1): same fixed-point iteration */
static inline double ll82_sigma_iter(double H, double d_minus_mu)
{
    double sigma = 1 / H;
    for (int it = 0; it < 10; ++it)
        sigma = 1 / H * SQRT_2_OVER_PI / (1 + erf(d_minus_mu / (sqrt(2.0) * sigma)));
    return sigma;
}

static inline hms ll82_f1(double dl, double dcoal)
{
    double dcoalCM = dcoal / CM, dlCM = dl / CM;
    hms r;
    r.H = 50.8 * pow(dlCM, -0.718);
    r.mu = dlCM;
    r.sigma = ll82_sigma_iter(r.H, dcoalCM - dlCM);
    return r;
}

static inline hms ll82_f2(double ds) /* :32-38 */
{
    double dsCM = ds / CM;
    hms r;
    r.H = 4.18 * pow(dsCM, -1.17);
    r.mu = dsCM;
    r.sigma = 1 / (sqrt(2 * PI) * r.H);
    return r;
}

static inline hms ll82_f3(double ds, double dl) /* :40-98 */
{
    double dsCM = ds / CM, dlCM = dl / CM;
    double Ff1 = dmax(0, (-2.25e4 * sq(dlCM - 0.403) - 37.9) * pow(dsCM, 2.5) +
                             9.67 * sq(dlCM - 0.170) + 4.95);
    double Ff2 = 1.02e4 * pow(dsCM, 2.83) + 2;
    double ds0 = dmax(0.04, pow(Ff1 / 2.83, 1 / 1.02e4));
    double Ff = dsCM > ds0 ? dmax(2.0, Ff1) : dmax(2.0, Ff2);
    double Dff3 = 0.241 * dsCM + 0.0129;
    double Pf301 = 1.68e5 * pow(dsCM, 2.33);
    double Pf302 = dmax(0, (43.4 * sq(dlCM + 1.81) - 159.0) / dsCM - 3870 * sq(dlCM - 0.285) - 58.1);
    double alpha = (dsCM - ds0) / (0.2 * ds0);
    double Pf303 = alpha * Pf301 + (1 - alpha) * Pf302;
    double Pf0 = dsCM < ds0 ? Pf301 : (dsCM > 1.2 * ds0 ? Pf302 : Pf303);
    hms r;
    r.sigma = 10 * Dff3;
    r.mu = log(Dff3) + sq(r.sigma);
    r.H = Pf0 * Dff3 / exp(-0.5 * sq(r.sigma));
    for (int it = 0; it < 10; ++it) {
        if (r.sigma == 0.0 || r.H == 0) {
            r.H = 0.0;
            r.mu = log(ds0);
            r.sigma = log(ds0);
            return r;
        }
        r.sigma = SQRT_2_OVER_PI * (Ff - 2) / r.H /
                  (1 - erf((log(0.01) - r.mu) / sqrt(2.0) / r.sigma));
        r.mu = log(Dff3) + sq(r.sigma);
        r.H = Pf0 * Dff3 / exp(-0.5 * sq(r.sigma));
    }
    return r;
}

static inline hms ll82_s1(double dl, double ds, double dcoal) /* :100-117 */
{
    double dsCM = ds / CM, dlCM = dl / CM, dcoalCM = dcoal / CM;
    hms r;
    r.H = 100 * exp(-3.25 * dsCM);
    r.mu = dlCM;
    r.sigma = ll82_sigma_iter(r.H, dcoalCM - dlCM);
    return r;
}

static inline hms ll82_s2(double dl, double ds, double St) /* :119-142 */
{
    double dsCM = ds / CM, dlCM = dl / CM;
    double Dss2 = 0.254 * pow(dsCM, 0.413) * exp(3.53 * pow(dsCM, 2.51) * (dlCM - dsCM));
    double bstar = 14.2 * exp(-17.2 * dsCM);
    double Ps20 = 0.23 * pow(dsCM, -3.93) * pow(dlCM, bstar);
    hms r;
    r.sigma = 10 * Dss2;
    r.mu = log(Dss2) + sq(r.sigma);
    r.H = Ps20 * Dss2 / exp(-0.5 * sq(r.sigma));
    double Fs = 5 * erf((St - 2.52e-6) / (1.85e-6)) + 6;
    for (int it = 0; it < 10; ++it) {
        r.sigma = SQRT_2_OVER_PI * (Fs - 1) / r.H /
                  (1 - erf((log(0.01) - r.mu) / sqrt(2.0) / r.sigma));
        r.mu = log(Dss2) + sq(r.sigma);
        r.H = Ps20 * Dss2 / exp(-0.5 * sq(r.sigma));
    }
    return r;
}

static inline hms ll82_d1(double W1, double dl, double dcoal, double CKE) /* :144-158 */
{
    double dlCM = dl / CM, dcoalCM = dcoal / CM;
    hms r;
    r.mu = dlCM * (1 - exp(-3.70 * (3.10 - W1)));
    r.H = 1.58e-5 * pow(CKE, -1.22);
    r.sigma = ll82_sigma_iter(r.H, dcoalCM - r.mu);
    return r;
}

static inline hms ll82_d2(double ds, double dl, double CKE) /* :160-190 */
{
    double dsCM = ds / CM, dlCM = dl / CM;
    double Ddd2 = exp(-17.4 * dsCM - 0.671 * (dlCM - dsCM)) * dsCM;
    double bstar = 0.007 * pow(dsCM, -2.54);
    double Pd20 = 0.0884 * pow(dsCM, -2.52) * pow(dlCM - dsCM, bstar);
    hms r, none;
    none.H = 0.0;
    none.mu = log(Ddd2);
    none.sigma = log(Ddd2);
    r.sigma = 10 * Ddd2;
    r.mu = log(Ddd2) + sq(r.sigma);
    r.H = Pd20 * Ddd2 / exp(-0.5 * sq(r.sigma));
    double Fd = dmax(1.0, 297.5 + 23.7 * log(CKE));
    if (Fd == 1.0) return none;
    for (int it = 0; it < 10; ++it) {
        if (r.sigma == 0.0 || r.H <= 0.1) return none;
        if (r.sigma >= 1.0) return none;
        r.sigma = SQRT_2_OVER_PI * (Fd - 1) / r.H /
                  (1 - erf((log(0.01) - r.mu) / sqrt(2.0) / r.sigma));
        r.mu = log(Ddd2) + sq(r.sigma);
        r.H = Pd20 * Ddd2 / exp(-0.5 * sq(r.sigma));
    }
    return r;
}

static inline double ll82_normal(hms m, double X) /* mu + sqrt(2) sigma erfinv(2X - 1) */
{
    return m.mu + sqrt(2.0) * m.sigma * erfinv_approx(2 * X - 1);
}

/* rand is updated in place, as in the reference */
ORACLE_API void oracle_ll82_fragmentation(const double *CKE, const double *W, const double *W2,
                                          const double *St, const double *ds, const double *dl,
                                          const double *dcoal, double *frag_volume, double *rand,
                                          double *Rf, double *Rs, double *Rd, i64 n, double tol)
{
    for (i64 i = 0; i < n; ++i) {
        if (dl[i] <= 0.4e-3) {
            frag_volume[i] = cube(dcoal[i]) * PI / 6;
            continue;
        }
        if (ds[i] == 0.0 || dl[i] == 0.0) {
            frag_volume[i] = 1e-18;
            continue;
        }
        /* ll82_Nr */
        Rf[i] = CKE[i] >= 0.893e-6 ? 1.11e-4 * pow(CKE[i], -0.654) : 1.0;
        /* (ll82_Nr is a conf.JIT_FLAGS fast-math helper: constants distributed + FMA, as in
         * straub_Nr above) */
        Rs[i] = W[i] >= 0.86 ? fma(-0.685, exp(fma(W2[i], -1.63, 1.63 * 0.86)), 0.685) : 0.0;
        Rd[i] = (Rs[i] + Rf[i]) > 1.0 ? 0.0 : 1.0 - (Rs[i] + Rf[i]);
        double fv;
        if (rand[i] <= Rf[i]) { /* filament */
            hms m1 = ll82_f1(dl[i], dcoal[i]), m2 = ll82_f2(ds[i]), m3 = ll82_f3(ds[i], dl[i]);
            double H1 = m1.H * m1.mu, H2 = m2.H * m2.mu, H3 = m3.H * exp(m3.mu);
            double Hsum = H1 + H2 + H3;
            rand[i] = rand[i] / Rf[i];
            if (rand[i] <= H1 / Hsum) {
                fv = ll82_normal(m1, dmax(rand[i] * Hsum / H1, tol));
            } else if (rand[i] <= (H1 + H2) / Hsum) {
                fv = ll82_normal(m2, (rand[i] * Hsum - H1) / H2);
            } else {
                fv = exp(ll82_normal(m3, dmin((rand[i] * Hsum - H1 - H2) / H3, 1.0 - tol)));
            }
        } else if (rand[i] <= Rf[i] + Rs[i]) { /* sheet */
            hms m1 = ll82_s1(dl[i], ds[i], dcoal[i]), m2 = ll82_s2(dl[i], ds[i], St[i]);
            double H1 = m1.H * m1.mu, H2 = m2.H * exp(m2.mu);
            double Hsum = H1 + H2;
            rand[i] = (rand[i] - Rf[i]) / (Rs[i]);
            if (rand[i] <= H1 / Hsum)
                fv = ll82_normal(m1, dmax(rand[i] * Hsum / H1, tol));
            else
                fv = exp(ll82_normal(m2, dmin((rand[i] * Hsum - H1) / H2, 1.0 - tol)));
        } else { /* disk */
            hms m1 = ll82_d1(W[i], dl[i], dcoal[i], CKE[i]), m2 = ll82_d2(ds[i], dl[i], CKE[i]);
            double H1 = m1.H * m1.mu, H2 = m2.H;
            double Hsum = H1 + H2;
            rand[i] = (rand[i] - Rf[i] - Rs[i]) / Rd[i];
            if (rand[i] <= H1 / Hsum)
                fv = ll82_normal(m1, dmax(rand[i] * Hsum / H1, tol));
            else
                fv = exp(ll82_normal(m2, dmin((rand[i] * Hsum - H1) / H2, 1 - tol)));
        }
        fv = fv * 0.01; /* cm -> m */
        frag_volume[i] = cube(fv) * PI / 6;
    }
}
