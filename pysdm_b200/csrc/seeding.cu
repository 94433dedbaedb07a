// seeding.cu -- particle injection into empty super-droplet slots.
// Reference: PySDM/backends/impl_numba/methods/seeding_methods.py:12-65 (serial, parallel=False):
// walk ALL n_sd slots in physical order; each slot with multiplicity 0, until
// n_inject have been filled, gets idx[i] = -1 and the multiplicity / extensive attributes of
// seed seeded_particle_index[k], k counting the filled slots.  Caller: particulator.py:447-490
// (which then resets idx and sanitizes), dynamic: dynamics/seeding.py:91-117.
//
// Parallel form: k = exclusive prefix count of empty slots before i (device-wide scan); slot
// i is filled iff it is empty and k < n_inject.  Same result as the serial walk, bit for bit;
// HBM-bound (8 B/SD read for the flags, 4 B/SD each way for the scan, scattered writes only
// for the injected slots).
#include "pysdm_b200.h"
#include "primitives.cuh"

namespace {

inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

__global__ void k_seeding_empty_flags(const i64 *__restrict__ multiplicity, i64 n_sd,
                                      u32 *__restrict__ empty)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_sd) empty[i] = multiplicity[i] == 0 ? 1u : 0u;
}

// empty_ps: exclusive prefix sums of the flags, empty_ps[n_sd] = number of empty slots
__global__ void k_seeding_fill(const u32 *__restrict__ empty_ps, i64 n_sd, i64 n_inject,
                               i64 *__restrict__ idx, i64 *__restrict__ multiplicity,
                               double *__restrict__ attributes, i64 n_attr,
                               const i64 *__restrict__ seed_index,
                               const i64 *__restrict__ seed_multiplicity,
                               const double *__restrict__ seed_attributes, i64 n_seeds,
                               i64 *__restrict__ n_injected)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_sd) return;
    if (i == 0) {
        const i64 total = (i64)empty_ps[n_sd];
        *n_injected = total < n_inject ? total : n_inject;
    }
    const i64 k = (i64)empty_ps[i];
    if (empty_ps[i + 1] == empty_ps[i] || k >= n_inject) return;
    idx[i] = -1;
    const i64 s = seed_index[k];
    multiplicity[i] = seed_multiplicity[s];
    for (i64 a = 0; a < n_attr; ++a) attributes[a * n_sd + i] = seed_attributes[a * n_seeds + s];
}

size_t seeding_ws_bytes(i64 n_sd)
{
    return align_up(sizeof(u32) * (size_t)(n_sd + 1)) +
           align_up(sizeof(u32) * scan_scratch_elems<u32>(n_sd + 1)) + align_up(sizeof(i64));
}

}  // namespace

PYSDM_EXPORT size_t pysdm_b200_seeding_workspace_bytes(i64 n_sd)
{
    return seeding_ws_bytes(n_sd < 1 ? 1 : n_sd);
}

// host_n_injected: host pointer; the stream is synchronised (the reference asserts the count)
PYSDM_EXPORT int pysdm_b200_seeding(i64 *idx, i64 *multiplicity, double *extensive_attributes,
                                    i64 n_attr, i64 n_sd, const i64 *seeded_particle_index,
                                    const i64 *seeded_particle_multiplicity,
                                    const double *seeded_particle_extensive_attributes,
                                    i64 n_seeds, i64 n_inject, i64 *host_n_injected, void *ws,
                                    size_t ws_bytes, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    *host_n_injected = 0;
    if (n_sd <= 0 || n_inject <= 0) return 0;  // the serial walk stops before the first slot
    if (ws_bytes < seeding_ws_bytes(n_sd)) return (int)cudaErrorMemoryAllocation;
    char *w = (char *)ws;
    u32 *empty = (u32 *)w;
    w += align_up(sizeof(u32) * (size_t)(n_sd + 1));
    u32 *scan_ws = (u32 *)w;
    w += align_up(sizeof(u32) * scan_scratch_elems<u32>(n_sd + 1));
    i64 *dev_count = (i64 *)w;
    const int g = grid_for(n_sd);
    k_seeding_empty_flags<<<g, kThreads, 0, st>>>(multiplicity, n_sd, empty);
    device_exclusive_scan<u32>(empty, empty, n_sd, 1, scan_ws, st);
    k_seeding_fill<<<g, kThreads, 0, st>>>(empty, n_sd, n_inject, idx, multiplicity,
                                           extensive_attributes, n_attr, seeded_particle_index,
                                           seeded_particle_multiplicity,
                                           seeded_particle_extensive_attributes, n_seeds,
                                           dev_count);
    int err = LAUNCH_CHECK();
    if (err) return err;
    err = (int)cudaMemcpyAsync(host_n_injected, dev_count, sizeof(i64), cudaMemcpyDeviceToHost, st);
    if (err) return err;
    return (int)cudaStreamSynchronize(st);
}
