// migration.cu -- the one exchange step of a multi-GPU run: after
// Displacement (PySDM/dynamics/displacement.py:103-153) super-droplets whose cell now belongs to
// another rank's slab of the grid are packed into records, exchanged (NCCL all-to-all-v,
// pysdm_b200/migration.py) and unpacked into free slots of the receiving rank.  The reference
// is single-process and has no counterpart; the kernels follow its data model: structure of
// arrays indexed by SD number, a permutation `idx` with a working length, `cell origin` /
// `position in cell` as [n_dims, n_sd] arrays (PySDM/impl/particle_attributes.py,
// PySDM/initialisation/..., environments/kinematic_2d.py:63-64).
//
// Slabs: the first grid dimension (nx rows) is split into `world` balanced contiguous ranges
// (pysdm_b200.sharding.cell_range); a rank's local cell ids are the global row-major ids minus
// the id of its first cell.
//
// A record is R = 1 + n_attr + 2 n_dims 8-byte words: multiplicity, the extensive attributes,
// the cell origin (global grid coordinates), the position in cell.  Pack and unpack run one
// thread per (record, word), so the record buffer is written / read fully coalesced and the
// SoA side is a gather / scatter of 8-byte words.
#include "pysdm_b200.h"

#include "common.cuh"

namespace {

struct Grid3 {
    i64 n[3];       // grid size per dimension (unused dimensions: 1)
    i64 stride[3];  // row-major strides of the global cell id
};

__device__ __forceinline__ i64 owner_of_row(i64 x, i64 base, i64 extra)
{
    const i64 split = extra * (base + 1);  // rows below `split` live on the `extra` larger slabs
    if (x < split) return x / (base + 1);
    return extra + (base > 0 ? (x - split) / base : 0);
}

// dest[i] = rank owning SD i's cell; cell_id[i] = its cell id local to THAT rank's slab
__global__ void k_destination(i64 *__restrict__ dest, i64 *__restrict__ cell_id,
                              const i64 *__restrict__ cell_origin, const i64 *__restrict__ idx,
                              i64 length, i64 n_sd, int n_dims, Grid3 g, i64 base, i64 extra)
{
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= length) return;
    const i64 i = idx[p];
    const i64 x = cell_origin[i];
    const i64 r = owner_of_row(x, base, extra);
    const i64 row_begin = r * base + (r < extra ? r : extra);
    i64 cid = (x - row_begin) * g.stride[0];
    for (int d = 1; d < n_dims; ++d) cid += cell_origin[(i64)d * n_sd + i] * g.stride[d];
    dest[i] = r;
    cell_id[i] = cid;
}

// records[r, :] <- SD order[q(r)], q skipping the stayers' segment [own_begin, own_end)
__global__ void k_pack(i64 *__restrict__ records, i64 n_rows, int n_words,
                       const i64 *__restrict__ order, i64 own_begin, i64 own_count,
                       i64 *__restrict__ multiplicity, const double *__restrict__ attributes,
                       int n_attr, const i64 *__restrict__ cell_origin,
                       const double *__restrict__ position_in_cell, int n_dims, i64 n_sd)
{
    const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_rows * n_words) return;
    const i64 r = t / n_words;
    const int w = (int)(t - r * n_words);
    const i64 i = order[r < own_begin ? r : r + own_count];
    i64 v;
    if (w == 0) {
        v = multiplicity[i];
        multiplicity[i] = 0;  // the slot is vacant from now on
    } else if (w < 1 + n_attr) {
        v = __double_as_longlong(attributes[(i64)(w - 1) * n_sd + i]);
    } else if (w < 1 + n_attr + n_dims) {
        v = cell_origin[(i64)(w - 1 - n_attr) * n_sd + i];
    } else {
        v = __double_as_longlong(position_in_cell[(i64)(w - 1 - n_attr - n_dims) * n_sd + i]);
    }
    records[t] = v;
}

// SD slots[r] <- records[r, :]; idx_tail[r] = slots[r]
__global__ void k_unpack(const i64 *__restrict__ records, i64 n_rows, int n_words,
                         const i64 *__restrict__ slots, i64 *__restrict__ idx_tail,
                         i64 *__restrict__ multiplicity, double *__restrict__ attributes,
                         int n_attr, i64 *__restrict__ cell_origin,
                         double *__restrict__ position_in_cell, int n_dims, i64 n_sd)
{
    const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_rows * n_words) return;
    const i64 r = t / n_words;
    const int w = (int)(t - r * n_words);
    const i64 i = slots[r];
    const i64 v = records[t];
    if (w == 0) {
        multiplicity[i] = v;
        idx_tail[r] = i;
    } else if (w < 1 + n_attr) {
        attributes[(i64)(w - 1) * n_sd + i] = __longlong_as_double(v);
    } else if (w < 1 + n_attr + n_dims) {
        cell_origin[(i64)(w - 1 - n_attr) * n_sd + i] = v;
    } else {
        position_in_cell[(i64)(w - 1 - n_attr - n_dims) * n_sd + i] = __longlong_as_double(v);
    }
}

}  // namespace

PYSDM_EXPORT int pysdm_b200_migration_destination(i64 *dest, i64 *cell_id, const i64 *cell_origin,
                                                  const i64 *idx, i64 length, i64 n_sd, int n_dims,
                                                  const i64 *host_grid, i64 world, void *stream)
{
    if (n_dims < 1 || n_dims > 3 || world < 1) return (int)cudaErrorInvalidValue;
    if (length <= 0) return 0;
    Grid3 g;
    for (int d = 0; d < 3; ++d) g.n[d] = d < n_dims ? host_grid[d] : 1;
    for (int d = 0; d < 3; ++d) g.stride[d] = 1;
    for (int d = n_dims - 2; d >= 0; --d) g.stride[d] = g.stride[d + 1] * g.n[d + 1];
    k_destination<<<grid_for(length), kThreads, 0, (cudaStream_t)stream>>>(
        dest, cell_id, cell_origin, idx, length, n_sd, n_dims, g, g.n[0] / world, g.n[0] % world);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_migration_pack(i64 *records, i64 n_rows, const i64 *order,
                                           i64 own_begin, i64 own_count, i64 *multiplicity,
                                           const double *attributes, i64 n_attr,
                                           const i64 *cell_origin, const double *position_in_cell,
                                           int n_dims, i64 n_sd, void *stream)
{
    if (n_rows <= 0) return 0;
    const int n_words = (int)(1 + n_attr + 2 * n_dims);
    k_pack<<<grid_for(n_rows * n_words), kThreads, 0, (cudaStream_t)stream>>>(
        records, n_rows, n_words, order, own_begin, own_count, multiplicity, attributes, (int)n_attr,
        cell_origin, position_in_cell, n_dims, n_sd);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_migration_unpack(const i64 *records, i64 n_rows, const i64 *slots,
                                             i64 *idx_tail, i64 *multiplicity, double *attributes,
                                             i64 n_attr, i64 *cell_origin, double *position_in_cell,
                                             int n_dims, i64 n_sd, void *stream)
{
    if (n_rows <= 0) return 0;
    const int n_words = (int)(1 + n_attr + 2 * n_dims);
    k_unpack<<<grid_for(n_rows * n_words), kThreads, 0, (cudaStream_t)stream>>>(
        records, n_rows, n_words, slots, idx_tail, multiplicity, attributes, (int)n_attr, cell_origin,
        position_in_cell, n_dims, n_sd);
    return LAUNCH_CHECK();
}
