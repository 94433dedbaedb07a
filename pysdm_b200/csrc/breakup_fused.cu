// breakup_fused.cu -- the breakup half of a Collision.step() in one launch
// (pysdm_b200_fused_breakup, include/pysdm_b200.h): for every pair slot with gamma != 0 the
// coalescence efficiency, the breakup efficiency, the fragment mass -- each a chain of pairwise
// Storage operations in the reference, evaluated here in registers in the same order -- the two
// uniforms, and collision_coalescence_breakup itself.
//
// Reference (paths relative to PySDM/):
//   dynamics/collisions/collision.py:206-234                        what a step with breakup issues
//   dynamics/collisions/coalescence_efficiencies/straub2010.py:24-47, constEc.py,
//     _parameterized.py:16-24 (Berry1967)
//   dynamics/collisions/breakup_efficiencies/constEb.py
//   dynamics/collisions/breakup_fragmentations/straub2010.py:44-102, always_n.py:12-15,
//     impl/volume_based.py:15-17
//   backends/impl_numba/methods/fragmentation_methods.py:218-260 (straub_fragmentation)
//   backends/impl_numba/methods/collisions_methods.py:246-311 (collision_coalescence_breakup)
//
// Pairwise operands follow pair_methods.py: op(data[idx[i]], data[idx[i + 1]]) with the pair
// already sorted by multiplicity (j first).  Only colliding pairs (a few per cent) do any of
// this; the reference evaluates the chains for every slot, but nothing else ever reads them.
#include "pysdm_b200.h"

#include "berry.cuh"
#include "breakup.cuh"
#include "dependents.cuh"
#include "frag.cuh"
#include "philox.cuh"

typedef pysdm_b200_breakup_args BreakupArgs;

namespace {

// Storage.divide_if_not_zero (impl_numba/storage_impl.py)
__device__ __forceinline__ double div_if_not_zero(double a, double b) { return b != 0.0 ? a / b : a; }

// element i of Random(seed) call `call` (philox.cuh: block i / 2, first or second double)
__device__ __forceinline__ double philox_element(u64 seed, u64 call, i64 i)
{
    u32 w[4];
    const u64 b = (u64)i >> 1;
    philox4x32_10((u32)b, (u32)(b >> 32), (u32)call, (u32)(call >> 32), (u32)seed,
                  (u32)(seed >> 32), w);
    return (i & 1) ? u01_from_words(w[2], w[3]) : u01_from_words(w[0], w[1]);
}

template <bool HandleAll>
__global__ void k_fused_breakup(const BreakupArgs a)
{
    const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    i64 cid = 0;
    PairOutcome o = {false, false, 0, 0, 0};
    if (i < a.length / 2) {
        const double g = a.gamma[i];
        if (g != 0.0) {  // pair_indices, collisions_methods.py:16-35
            const i64 offset = 1 - (i64)a.is_first_in_pair[2 * i];
            const i64 j = a.idx[2 * i + offset], k = a.idx[2 * i + 1 + offset];
            const double u_proc = a.rand_proc ? a.rand_proc[i] : philox_element(a.seed, a.call_proc, i);
            const double *mass = a.attributes + a.mass_row * a.n_sd;
            double v_j = 0, v_k = 0, r_j = 0, r_k = 0, w_j = 0, w_k = 0;
            if (a.ec_kind == 1 || a.nf_kind == 1) {
                v_j = a.volume[j];
                v_k = a.volume[k];
                w_j = a.fall_velocity[j];
                w_k = a.fall_velocity[k];
            }
            if (a.ec_kind == 2 || a.nf_kind == 1) {
                r_j = a.radius[j];
                r_k = a.radius[k];
            }
            // ---- coalescence efficiency --------------------------------------------------
            double ec;
            if (a.ec_kind == 0) {
                ec = a.ec_const;
            } else if (a.ec_kind == 1) {  // straub2010.py:24-47
                double tmp = v_j + v_k;                 // tmp.sum(volume)
                double Sc = tmp * a.six_over_pi;        // Sc.fill(tmp); Sc *= 6 / PI
                tmp = tmp * 2;                          // tmp *= 2
                double d = fabs(w_j - w_k);             // tmp2.distance(fall velocity)
                const double tmp2 = signed_pow(d, 2.0); // tmp2 **= 2
                double We = v_j * v_k;                  // We.multiply(volume)
                We = div_if_not_zero(We, tmp);
                We = We * tmp2;
                We = We * a.rho_w;
                Sc = signed_pow(Sc, a.two_thirds);      // Sc **= 2 / 3
                Sc = Sc * a.pi_sgm_w;                   // Sc *= PI * sgm_w
                We = div_if_not_zero(We, Sc);
                We = We * -1.15;
                ec = exp(We);                           // output.fill(np.exp(We))
            } else {  // _parameterized.py:16-24: linear_collection_efficiency; output **= 2
                Berry q;
#pragma unroll
                for (int e = 0; e < 13; ++e) q.p[e] = a.ec_params[e];
                q.int_mask = a.ec_int_mask;
                ec = signed_pow(berry_efficiency(q, r_j, r_k, a.ec_unit), 2.0);
            }
            // ---- fragment mass (only pairs that neither bounce nor coalesce need it) ------
            const double eb = a.eb_const;
            double fragment_mass = 0.0;
            const bool bouncing = u_proc - (ec + (1 - ec) * eb) > 0;
            if (!bouncing && !(u_proc - ec < 0)) {
                if (a.nf_kind == 0) {  // always_n.py:12-15: frag_mass.sum(water mass); /= N
                    fragment_mass = (mass[j] + mass[k]) / a.nf_n;
                } else {  // straub2010.py:44-102
                    const double u_frag =
                        a.rand_frag ? a.rand_frag[i] : philox_element(a.seed, a.call_frag, i);
                    const double max_size = v_j > v_k ? v_j : v_k;       // max_size.max(volume)
                    const double x_plus_y = v_j + v_k;                  // sum_of_volumes
                    double ds = (r_k < r_j ? r_k : r_j) * 2;            // ds.min(radius); *= 2
                    const double tmp = v_j + v_k;
                    double Sc = signed_pow(tmp, a.two_thirds);          // Sc.fill(tmp); **= 2 / 3
                    Sc = Sc * a.nf_sc_factor;
                    double d = fabs(w_j - w_k);
                    const double tmp2 = signed_pow(d, 2.0);
                    double CKE = v_j * v_k;                             // CKE.multiply(volume)
                    CKE = div_if_not_zero(CKE, tmp);
                    CKE = CKE * tmp2;
                    CKE = CKE * a.half_rho_w;
                    double We = div_if_not_zero(CKE, Sc);               // We.fill(CKE); / Sc
                    double CW = We * CKE;                               // CW.fill(We); *= CKE
                    CW = CW / a.micro_joule;                            // CW /= si.uJ
                    double gam = r_j > r_k ? r_j : r_k;                 // gam.max(radius)
                    const double rmin = r_k < r_j ? r_k : r_j;          // tmp.min(radius)
                    gam = div_if_not_zero(gam, rmin);
                    Limits lim;
                    lim.vmin = a.vmin;
                    lim.nfmax = a.nfmax;
                    lim.has_nfmax = a.has_nfmax;
                    StraubResult r;
                    r.Nr1 = r.Nr2 = r.Nr3 = 0.0;  // straub_tmp[...].fill(0)
                    straub_fragment(CW, gam, ds, max_size, x_plus_y, u_frag, lim, r);
                    fragment_mass = a.rho_w * r.frag_volume;  // mass_of_water_volume
                }
            }
            o = collide_pair<HandleAll>(j, k, g, u_proc, ec, eb, fragment_mass, a.multiplicity,
                                        a.attributes, a.n_attr, a.n_sd, mass, a.max_multiplicity,
                                        a.healthy);
            if (o.active) {
                cid = a.cell_id[j];
                if (a.dep_radius != nullptr) {  // derived attributes of the two changed SDs
                    DependentParams dp;
                    dp.rho_w = a.rho_w;
                    dp.radius_scale = a.dep_radius_scale;
                    dp.radius_exponent = a.dep_radius_exponent;
                    dp.table_factor = a.dep_table_factor;
                    dp.radius_max = a.dep_radius_max;
                    dp.table_b = a.dep_table_b;
                    dp.table_c = a.dep_table_c;
                    const i64 sd[2] = {j, k};
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        double vol, rad, vel;
                        const bool ok = dependents_of(mass[sd[e]], dp, vol, rad, vel);
                        if (!ok && a.dep_status != nullptr) a.dep_status[0] = 1;
                        if (a.dep_volume != nullptr) a.dep_volume[sd[e]] = vol;
                        a.dep_radius[sd[e]] = rad;
                        if (ok && a.dep_velocity != nullptr) a.dep_velocity[sd[e]] = vel;
                    }
                }
            }
        }
    }
    cell_counter_add(a.coalescence_rate, cid, o.coal_add, o.active);
    cell_counter_add(a.breakup_rate, cid, o.break_add, o.active);
    cell_counter_add(a.breakup_rate_deficit, cid, o.deficit_add, o.active);
    if (o.overflow && a.n_overflow) atomicAdd((unsigned long long *)a.n_overflow, 1ull);
}

}  // namespace

PYSDM_EXPORT int pysdm_b200_fused_breakup(const pysdm_b200_breakup_args *host_args, void *stream)
{
    const BreakupArgs a = *host_args;
    if (a.n_sd <= 0 || a.length < 0 || a.length > a.n_sd) return (int)cudaErrorInvalidValue;
    if (a.ec_kind < 0 || a.ec_kind > 2 || a.nf_kind < 0 || a.nf_kind > 1)
        return (int)cudaErrorInvalidValue;
    if ((a.ec_kind == 1 || a.nf_kind == 1) && (!a.volume || !a.fall_velocity))
        return (int)cudaErrorInvalidValue;
    if ((a.ec_kind == 2 || a.nf_kind == 1) && !a.radius) return (int)cudaErrorInvalidValue;
    if ((a.rand_proc == nullptr) != (a.rand_frag == nullptr)) return (int)cudaErrorInvalidValue;
    const i64 n_pairs = a.length / 2;
    if (n_pairs <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (a.handle_all_breakups)
        k_fused_breakup<true><<<grid_for(n_pairs), kThreads, 0, st>>>(a);
    else
        k_fused_breakup<false><<<grid_for(n_pairs), kThreads, 0, st>>>(a);
    return LAUNCH_CHECK();
}
