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

#include "primitives.cuh"

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

// =====================================================================================
// Neighbour exchange without host round trips (pysdm_b200.migration.SlabMigrator.migrate):
// with slabs of rows and Courant numbers below one, super-droplets only move to the two ring
// neighbours.  Everything that depends on the counts stays on the device:
//   classify   dest / local cell id per SD, class (0 stay, 1 to rank+1, 2 to rank-1) packed as
//              (up << 32 | down) for ONE 64-bit exclusive scan -> stable positions of all three
//   split      stayers -> idx_out[0 .. n_stay), leavers -> fixed-capacity record buffers
//              (row 0 = header with the count), their slots onto the device free stack
//   (exchange  two send/recv pairs of fixed size: NCCL, no count exchange)
//   merge      arrivals (source rank order) -> slots popped from the free stack, appended to
//              idx_out, local cell ids; counters updated
// counters (i64): [0] n_stay  [1] n_free  [2] length out  [3] status bits (1: some droplets
// moved farther than a neighbour slab -- they stay in idx for the general path to move; 2: a
// buffer or the SD arrays ran out of capacity -- fatal, nothing more is moved)
//            [4] n_up  [5] n_down  [6] arrivals from below  [7] arrivals from above
// =====================================================================================
namespace {

__global__ void k_mig_classify(const i64 *__restrict__ idx, i64 length,
                               const i64 *__restrict__ cell_origin, i64 n_sd, int n_dims, Grid3 g,
                               i64 base, i64 extra, i64 rank, i64 world, i64 *__restrict__ dest,
                               i64 *__restrict__ cell_id, u64 *__restrict__ cls,
                               i64 *__restrict__ counters)
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
    const i64 up = (rank + 1) % world, down = (rank + world - 1) % world;
    u64 c = 0;
    if (r != rank) {
        if (r == up)
            c = (u64)1 << 32;
        else if (r == down)
            c = 1;
        else
            atomicOr((unsigned long long *)&counters[3], 1ull);  // left to the general path
    }
    cls[p] = c;
}

// one thread per position: stayers keep their order, leavers go to their buffer in order
__global__ void k_mig_split(const i64 *__restrict__ idx, i64 length, const u64 *__restrict__ cls,
                            const u64 *__restrict__ pos /* exclusive scan of cls, [length + 1] */,
                            i64 *__restrict__ idx_out, i64 *__restrict__ leavers /* [2, cap] */,
                            i64 cap, i64 *__restrict__ counters)
{
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p > length) return;
    const u64 before = pos[p];
    const i64 up_before = (i64)(before >> 32), down_before = (i64)(before & 0xffffffffu);
    if (p == length) {  // totals
        counters[0] = length - up_before - down_before;
        counters[4] = up_before;
        counters[5] = down_before;
        if (up_before > cap || down_before > cap)
            atomicOr((unsigned long long *)&counters[3], 2ull);
        return;
    }
    const u64 c = cls[p];
    const i64 i = idx[p];
    if (c == 0) {
        idx_out[p - up_before - down_before] = i;
    } else if (c >> 32) {
        if (up_before < cap) leavers[up_before] = i;
    } else {
        if (down_before < cap) leavers[cap + down_before] = i;
    }
}

// records of both directions: buffer d = [cap + 1, n_words], row 0 = header (word 0 = count)
__global__ void k_mig_pack2(i64 *__restrict__ send /* [2, cap + 1, n_words] */, i64 cap,
                            int n_words, const i64 *__restrict__ leavers,
                            i64 *__restrict__ multiplicity, const double *__restrict__ attributes,
                            int n_attr, const i64 *__restrict__ cell_origin,
                            const double *__restrict__ position_in_cell, int n_dims, i64 n_sd,
                            i64 *__restrict__ free_stack, i64 *__restrict__ counters)
{
    const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2 * cap * n_words) return;
    const i64 status = counters[3] & 2;
    const int d = (int)(t / (cap * n_words));
    const i64 rem = t - (i64)d * cap * n_words;
    const i64 r = rem / n_words;
    const int w = (int)(rem - r * n_words);
    const i64 count = status ? 0 : counters[4 + d];
    i64 *buf = send + (i64)d * (cap + 1) * n_words;
    if (r == 0 && w == 0) buf[0] = count;
    if (r >= count) return;
    const i64 i = leavers[(i64)d * cap + r];
    i64 v;
    if (w == 0) {
        v = multiplicity[i];
        multiplicity[i] = 0;  // the slot is vacant from now on
        free_stack[counters[1] + (d ? counters[4] : 0) + r] = i;
    } else if (w < 1 + n_attr) {
        v = __double_as_longlong(attributes[(i64)(w - 1) * n_sd + i]);
    } else if (w < 1 + n_attr + n_dims) {
        v = cell_origin[(i64)(w - 1 - n_attr) * n_sd + i];
    } else {
        v = __double_as_longlong(position_in_cell[(i64)(w - 1 - n_attr - n_dims) * n_sd + i]);
    }
    buf[(r + 1) * n_words + w] = v;
}

// arrivals: source s = 0 (lower of the two source ranks) then s = 1, appended behind the stayers
__global__ void k_mig_merge(const i64 *__restrict__ recv /* [2, cap + 1, n_words] */, i64 cap,
                            int n_words, i64 *__restrict__ idx_out,
                            i64 *__restrict__ multiplicity, double *__restrict__ attributes,
                            int n_attr, i64 *__restrict__ cell_origin,
                            double *__restrict__ position_in_cell, int n_dims, i64 n_sd,
                            const i64 *__restrict__ free_stack, i64 *__restrict__ counters)
{
    const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2 * cap * n_words) return;
    if (counters[3] & 2) return;
    const int s = (int)(t / (cap * n_words));
    const i64 rem = t - (i64)s * cap * n_words;
    const i64 r = rem / n_words;
    const int w = (int)(rem - r * n_words);
    const i64 *buf = recv + (i64)s * (cap + 1) * n_words;
    const i64 count0 = recv[0], count1 = recv[(cap + 1) * n_words];
    const i64 count = s ? count1 : count0;
    if (r >= count) return;
    const i64 a = (s ? count0 : 0) + r;  // arrival number
    const i64 n_free = counters[1] + counters[4] + counters[5];  // after the leavers' push
    if (count0 + count1 > n_free) {  // no room (capacity of the SD arrays)
        if (r == 0 && w == 0) atomicOr((unsigned long long *)&counters[3], 2ull);
        return;
    }
    const i64 i = free_stack[n_free - 1 - a];
    const i64 v = buf[(r + 1) * n_words + w];
    if (w == 0) {
        multiplicity[i] = v;
        idx_out[counters[0] + a] = i;
    } else if (w < 1 + n_attr) {
        attributes[(i64)(w - 1) * n_sd + i] = __longlong_as_double(v);
    } else if (w < 1 + n_attr + n_dims) {
        cell_origin[(i64)(w - 1 - n_attr) * n_sd + i] = v;
    } else {
        position_in_cell[(i64)(w - 1 - n_attr - n_dims) * n_sd + i] = __longlong_as_double(v);
    }
}

// local cell ids of the arrivals (one thread per arrival, from the origin words of its record)
__global__ void k_mig_cell_ids(const i64 *__restrict__ recv, i64 cap, int n_words, int n_attr,
                               int n_dims, Grid3 g, i64 row_begin, i64 *__restrict__ cell_id,
                               const i64 *__restrict__ free_stack,
                               const i64 *__restrict__ counters)
{
    const i64 a = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (counters[3] & 2) return;
    const i64 count0 = recv[0], total = count0 + recv[(cap + 1) * n_words];
    const i64 n_free = counters[1] + counters[4] + counters[5];
    if (a >= total || total > n_free) return;
    const i64 *rec = a < count0 ? recv + (a + 1) * n_words
                                : recv + ((cap + 1) + (a - count0) + 1) * n_words;
    i64 cid = (rec[1 + n_attr] - row_begin) * g.stride[0];
    for (int d = 1; d < n_dims; ++d) cid += rec[1 + n_attr + d] * g.stride[d];
    cell_id[free_stack[n_free - 1 - a]] = cid;
}

__global__ void k_mig_finish(const i64 *__restrict__ recv, i64 cap, int n_words,
                             i64 *__restrict__ counters)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (counters[3] & 2) return;
    const i64 c0 = recv[0], c1 = recv[(cap + 1) * n_words];
    counters[6] = c0;
    counters[7] = c1;
    counters[2] = counters[0] + c0 + c1;
    counters[1] = counters[1] + counters[4] + counters[5] - c0 - c1;
}

}  // namespace

PYSDM_EXPORT size_t pysdm_b200_migration_split_workspace_bytes(i64 length)
{
    return 2 * ((sizeof(u64) * (size_t)(length + 1) + 255) / 256 * 256) +
           (sizeof(u64) * scan_scratch_elems<u64>(length + 1) + 255) / 256 * 256;
}

// classify + scan + split + pack: leaves the stayers in idx_out[0 .. counters[0]) and the two
// send buffers filled; modifies nothing (status 1) when a droplet moved beyond a neighbour slab
PYSDM_EXPORT int pysdm_b200_migration_split(
    const i64 *idx, i64 length, i64 *idx_out, i64 *dest, i64 *cell_id, const i64 *cell_origin,
    const double *position_in_cell, i64 *multiplicity, const double *attributes, i64 n_attr,
    i64 n_sd, int n_dims, const i64 *host_grid, i64 rank, i64 world, i64 *send, i64 *leavers,
    i64 cap, i64 *free_stack, i64 *counters, void *ws, size_t ws_bytes, void *stream)
{
    if (n_dims < 1 || n_dims > 3 || world < 1 || cap < 1) return (int)cudaErrorInvalidValue;
    if (ws_bytes < pysdm_b200_migration_split_workspace_bytes(length))
        return (int)cudaErrorMemoryAllocation;
    cudaStream_t st = (cudaStream_t)stream;
    Grid3 g;
    for (int d = 0; d < 3; ++d) g.n[d] = d < n_dims ? host_grid[d] : 1;
    for (int d = 0; d < 3; ++d) g.stride[d] = 1;
    for (int d = n_dims - 2; d >= 0; --d) g.stride[d] = g.stride[d + 1] * g.n[d + 1];
    const size_t seg = (sizeof(u64) * (size_t)(length + 1) + 255) / 256 * 256;
    u64 *cls = (u64 *)ws, *pos = (u64 *)((char *)ws + seg), *scan_ws = (u64 *)((char *)ws + 2 * seg);
    const int n_words = (int)(1 + n_attr + 2 * n_dims);
    if (length > 0)
        k_mig_classify<<<grid_for(length), kThreads, 0, st>>>(
            idx, length, cell_origin, n_sd, n_dims, g, g.n[0] / world, g.n[0] % world, rank, world,
            dest, cell_id, cls, counters);
    device_exclusive_scan<u64>(cls, pos, length, 1, scan_ws, st);
    k_mig_split<<<grid_for(length + 1), kThreads, 0, st>>>(idx, length, cls, pos, idx_out, leavers,
                                                           cap, counters);
    k_mig_pack2<<<grid_for(2 * cap * n_words), kThreads, 0, st>>>(
        send, cap, n_words, leavers, multiplicity, attributes, (int)n_attr, cell_origin,
        position_in_cell, n_dims, n_sd, free_stack, counters);
    return LAUNCH_CHECK();
}

// after the exchange: arrivals into free slots, appended to idx_out; counters[2] = new length
PYSDM_EXPORT int pysdm_b200_migration_merge(
    const i64 *recv, i64 cap, i64 *idx_out, i64 *multiplicity, double *attributes, i64 n_attr,
    i64 *cell_origin, double *position_in_cell, i64 *cell_id, i64 n_sd, int n_dims,
    const i64 *host_grid, i64 rank, i64 world, const i64 *free_stack, i64 *counters, void *stream)
{
    if (n_dims < 1 || n_dims > 3 || world < 1 || cap < 1) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    Grid3 g;
    for (int d = 0; d < 3; ++d) g.n[d] = d < n_dims ? host_grid[d] : 1;
    for (int d = 0; d < 3; ++d) g.stride[d] = 1;
    for (int d = n_dims - 2; d >= 0; --d) g.stride[d] = g.stride[d + 1] * g.n[d + 1];
    const i64 base = g.n[0] / world, extra = g.n[0] % world;
    const i64 row_begin = rank * base + (rank < extra ? rank : extra);
    const int n_words = (int)(1 + n_attr + 2 * n_dims);
    k_mig_cell_ids<<<grid_for(2 * cap), kThreads, 0, st>>>(
        recv, cap, n_words, (int)n_attr, n_dims, g, row_begin, cell_id, free_stack, counters);
    k_mig_merge<<<grid_for(2 * cap * n_words), kThreads, 0, st>>>(
        recv, cap, n_words, idx_out, multiplicity, attributes, (int)n_attr, cell_origin,
        position_in_cell, n_dims, n_sd, free_stack, counters);
    k_mig_finish<<<1, 32, 0, st>>>(recv, cap, n_words, counters);
    return LAUNCH_CHECK();
}
