"""Migration of super-droplets between the slabs of a multi-GPU run -- the one step of the
path that needs communication: collisions pair droplets of one cell only, so
cells are sharded over ranks as slabs of the first grid dimension and the collision step runs
without any collective; after `Displacement` (PySDM/dynamics/displacement.py:103-153) the
droplets whose cell now lies in another rank's slab change hands.

    destination  (CUDA)  dest[i] = owner rank, cell_id[i] = cell id local to that rank
    sort         (CUDA)  stable counting sort of idx by destination -> per-rank segments + counts
    pack         (CUDA)  leavers -> records [n_send, R] grouped by destination
    exchange     (NCCL)  counts, then records: torch.distributed.all_to_all_single (all-to-all-v)
    unpack       (CUDA)  arrivals -> free SD slots, appended to idx

With slabs of rows and Courant numbers below one droplets only ever move to the two ring
neighbours; `migrate()` therefore first runs a device-driven neighbour exchange with NO host
round trip before the data is where it belongs (csrc/migration.cu, second half): one 64-bit
scan gives the stable positions of stayers / up-movers / down-movers, the leavers are packed
into two fixed-capacity buffers whose first row carries the count, two send/recv pairs move
them (NCCL, fixed sizes: no count exchange), the arrivals are merged from a device free stack;
the host reads the new working length once, at the end.  Droplets that moved farther (thin
slabs) are left in place by that pass and handled by the general all-to-all-v path above,
which all ranks enter together when any rank saw such a droplet (one all-reduced flag).

The reference is single-process (no counterpart to cite); the data model is its own
(PySDM/impl/particle_attributes.py: SoA by SD number + `idx` + working length).  `exchange`
is plain torch.distributed and runs with gloo on CPU tensors too, which is how the routing
logic is tested without GPUs (tests/test_sharding.py); pack/unpack have no CPU path.
"""

import ctypes
import os

import torch

from . import _lib
from . import methods as m
from .sharding import cell_range


def exchange(send, send_counts, group=None):
    """all-to-all-v of record rows.  `send`: [n_send, R] tensor whose rows are grouped by
    destination rank in rank order, `send_counts[r]` rows for rank r.  Returns
    (recv [n_recv, R], recv_counts list) -- rows grouped by source rank in rank order."""
    import torch.distributed as dist  # pylint: disable=import-outside-toplevel

    world = dist.get_world_size(group)
    assert len(send_counts) == world and sum(send_counts) == send.shape[0]
    counts_out = torch.tensor(send_counts, dtype=torch.int64, device=send.device)
    counts_in = torch.empty_like(counts_out)
    dist.all_to_all_single(counts_in, counts_out, group=group)
    recv_counts = [int(c) for c in counts_in.tolist()]
    recv = torch.empty((sum(recv_counts), send.shape[1]), dtype=send.dtype, device=send.device)
    dist.all_to_all_single(recv, send, output_split_sizes=recv_counts,
                           input_split_sizes=list(send_counts), group=group)
    return recv, recv_counts


class SlabMigrator:  # pylint: disable=too-many-instance-attributes
    """Owns nothing but the bookkeeping: the SD arrays are the caller's (device tensors with
    capacity n_cap >= the largest population the rank will ever hold):
    multiplicity i64[n_cap], attributes f64[A, n_cap], cell_origin i64[D, n_cap] (GLOBAL grid
    coordinates), position_in_cell f64[D, n_cap], cell_id i64[n_cap] (local to the slab),
    idx i64[n_cap] with working length `length`; slots not referenced by idx[:length] are
    free."""

    def __init__(self, *, grid, rank, world, multiplicity, attributes, cell_origin,
                 position_in_cell, cell_id, idx, length, exchange_fn=exchange,
                 neighbour_exchange_fn=None, capacity=None, neighbours_only=False):
        self.grid = tuple(int(n) for n in grid)
        self.rank, self.world = int(rank), int(world)
        self.rows = cell_range(self.grid[0], self.rank, self.world)
        self.multiplicity, self.attributes = multiplicity, attributes
        self.cell_origin, self.position_in_cell = cell_origin, position_in_cell
        self.cell_id, self.idx, self.length = cell_id, idx, int(length)
        self.n_cap = multiplicity.numel()
        self.n_dims, self.n_attr = cell_origin.shape[0], attributes.shape[0]
        self.n_words = 1 + self.n_attr + 2 * self.n_dims
        self.exchange_fn = exchange_fn
        dev = multiplicity.device
        self.dest = torch.zeros(self.n_cap, dtype=torch.int64, device=dev)
        self.order = torch.empty(self.n_cap, dtype=torch.int64, device=dev)
        self.rank_idx = torch.arange(self.world, dtype=torch.int64, device=dev)
        self.dest_start = torch.zeros(self.world + 1, dtype=torch.int64, device=dev)
        # free slots: everything idx[:length] does not reference; kept as a stack on the device
        used = torch.zeros(self.n_cap, dtype=torch.bool, device=dev)
        used[idx[:self.length]] = True
        self.free_stack = torch.zeros(self.n_cap, dtype=torch.int64, device=dev)
        self.n_free = 0
        self.free = torch.nonzero(~used).reshape(-1)
        self.n_sent = self.n_received = 0
        # neighbour exchange: records per direction, buffers, device counters (csrc/migration.cu)
        self.neighbour_exchange_fn = neighbour_exchange_fn or neighbour_exchange
        self.neighbours_only = bool(neighbours_only)
        # (a few per cent of the population per direction; exceeding it is reported, not silent)
        self.cap = int(capacity or os.environ.get("PYSDM_MIG_CAP") or max(4096, self.n_cap // 64))
        self.send = torch.zeros((2, self.cap + 1, self.n_words), dtype=torch.int64, device=dev)
        self.recv = torch.zeros_like(self.send)
        self.leavers = torch.zeros((2, self.cap), dtype=torch.int64, device=dev)
        self.idx_out = torch.zeros(self.n_cap, dtype=torch.int64, device=dev)
        self.counters = torch.zeros(8, dtype=torch.int64, device=dev)
        self.counters_host = torch.zeros(8, dtype=torch.int64).pin_memory()
        self.far = torch.zeros(1, dtype=torch.int64, device=dev)
        self._ws = None
        self.n_general = 0  # times the general path had to run

    @property
    def free(self):
        return self.free_stack[:self.n_free]

    @free.setter
    def free(self, slots):
        self.n_free = int(slots.numel())
        self.free_stack[:self.n_free] = slots

    @property
    def n_local_cells(self):
        n = self.rows[1] - self.rows[0]
        for g in self.grid[1:]:
            n *= g
        return n

    def _destination(self, idx, length):
        grid = (ctypes.c_int64 * self.n_dims)(*self.grid)
        _lib.check(_lib.load().pysdm_b200_migration_destination(
            self.dest.data_ptr(), self.cell_id.data_ptr(), self.cell_origin.data_ptr(),
            idx.data_ptr(), int(length), self.n_cap, self.n_dims, grid, self.world,
            torch.cuda.current_stream().cuda_stream), "migration_destination")

    def migrate(self):
        """one exchange; returns the new working length.  Afterwards idx[:length] holds the
        stayers in their previous relative order followed by the arrivals (by source rank),
        cell_id is valid (slab-local) for all of them, and the population is NOT sorted by cell."""
        self.split()
        self.neighbour_exchange_fn(self.send, self.recv, self.rank, self.world)
        if self.neighbours_only:
            # the caller vouches that no droplet moves farther than one slab per exchange
            # (Courant numbers below one with slabs at least one row thick): no agreement
            # between the ranks is needed, so no collective couples them all in every step
            if self.merge(all_reduce_far=False):
                raise RuntimeError(f"rank {self.rank}: a droplet moved beyond a neighbour slab "
                                   "although the migrator was built with neighbours_only=True")
        elif self.merge(all_reduce_far=self.world > 1):  # some rank saw a far mover: all ranks
            self.migrate_general()                       # run the general path on what is left
        return self.length

    def migrate_general(self):
        """the all-to-all-v path: any destination, two host round trips"""
        self.n_general += 1
        records, send_counts = self.pack()
        recv, _ = self.exchange_fn(records, send_counts)
        return self.unpack(recv)

    # ---- neighbour path (no host round trip until the end) -----------------------------------
    def split(self):
        """classify + scan + split + pack: the stayers in idx_out, both send buffers filled"""
        lib, stream = _lib.load(), torch.cuda.current_stream().cuda_stream
        nbytes = lib.pysdm_b200_migration_split_workspace_bytes(self.n_cap)
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = torch.empty(int(nbytes) + 256, dtype=torch.uint8,
                                   device=self.multiplicity.device)
        self.counters.zero_()
        self.counters[1] = self.n_free
        grid = (ctypes.c_int64 * self.n_dims)(*self.grid)
        _lib.check(lib.pysdm_b200_migration_split(
            self.idx.data_ptr(), int(self.length), self.idx_out.data_ptr(), self.dest.data_ptr(),
            self.cell_id.data_ptr(), self.cell_origin.data_ptr(),
            self.position_in_cell.data_ptr(), self.multiplicity.data_ptr(),
            self.attributes.data_ptr(), self.n_attr, self.n_cap, self.n_dims, grid, self.rank,
            self.world, self.send.data_ptr(), self.leavers.data_ptr(), self.cap,
            self.free_stack.data_ptr(), self.counters.data_ptr(), self._ws.data_ptr(),
            self._ws.numel(), stream), "migration_split")

    def merge(self, all_reduce_far=False):
        """arrivals (self.recv) into free slots behind the stayers; reads the counters back
        (the one host synchronisation of the exchange).  Returns True when droplets that moved
        beyond a neighbour slab are still waiting for the general path (on any rank when
        `all_reduce_far`)"""
        lib, stream = _lib.load(), torch.cuda.current_stream().cuda_stream
        grid = (ctypes.c_int64 * self.n_dims)(*self.grid)
        _lib.check(lib.pysdm_b200_migration_merge(
            self.recv.data_ptr(), self.cap, self.idx_out.data_ptr(), self.multiplicity.data_ptr(),
            self.attributes.data_ptr(), self.n_attr, self.cell_origin.data_ptr(),
            self.position_in_cell.data_ptr(), self.cell_id.data_ptr(), self.n_cap, self.n_dims,
            grid, self.rank, self.world, self.free_stack.data_ptr(), self.counters.data_ptr(),
            stream), "migration_merge")
        if all_reduce_far:
            import torch.distributed as dist  # pylint: disable=import-outside-toplevel

            self.far.copy_(self.counters[3:4] & 1)
            dist.all_reduce(self.far, op=dist.ReduceOp.MAX)
            self.counters[3:4] |= self.far
        self.counters_host.copy_(self.counters, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        c = [int(x) for x in self.counters_host]
        if c[3] & 2:
            raise RuntimeError(f"rank {self.rank}: migration capacity exceeded (leavers {c[4]} + "
                               f"{c[5]} of {self.cap} per direction, or no free slot left of "
                               f"{self.n_cap})")
        self.length, self.n_free = c[2], c[1]
        self.idx[:self.length] = self.idx_out[:self.length]
        self.n_sent += c[4] + c[5]
        self.n_received += c[6] + c[7]
        return bool(c[3] & 1)

    def pack(self):
        """first half: (records [n_send, R] grouped by destination rank, send_counts)"""
        lib, stream = _lib.load(), torch.cuda.current_stream().cuda_stream
        start = [0] * (self.world + 1)
        if self.length > 0:
            self._destination(self.idx, self.length)
            # idx sorted by destination (stable): segment r = [dest_start[r], dest_start[r + 1])
            m.counting_sort_by_cell_id(self.order, self.idx, self.dest, self.rank_idx,
                                       self.length, self.dest_start)
            start = [int(x) for x in self.dest_start.tolist()]  # host sync: the counts
        own_begin, own_end = start[self.rank], start[self.rank + 1]
        own_count = own_end - own_begin
        send_counts = [start[r + 1] - start[r] if r != self.rank else 0
                       for r in range(self.world)]
        n_send = sum(send_counts)
        records = torch.empty((n_send, self.n_words), dtype=torch.int64,
                              device=self.multiplicity.device)
        _lib.check(lib.pysdm_b200_migration_pack(
            records.data_ptr(), n_send, self.order.data_ptr(), own_begin, own_count,
            self.multiplicity.data_ptr(), self.attributes.data_ptr(), self.n_attr,
            self.cell_origin.data_ptr(), self.position_in_cell.data_ptr(), self.n_dims,
            self.n_cap, stream), "migration_pack")
        # the leavers' slots become free; the stayers keep their order
        leavers = torch.cat((self.order[:own_begin], self.order[own_end:self.length]))
        self.idx[:own_count] = self.order[own_begin:own_end]
        self.free = torch.cat((self.free, leavers))
        self.length = own_count
        self.n_sent += n_send
        return records, send_counts

    def unpack(self, recv):
        """second half: arrivals [n_recv, R] into free slots, appended to idx"""
        n_recv = recv.shape[0]
        if n_recv > self.free.numel():
            raise RuntimeError(f"rank {self.rank}: {n_recv} arriving super-droplets, "
                               f"{self.free.numel()} free slots (capacity {self.n_cap})")
        if n_recv:
            slots = self.free[self.free.numel() - n_recv:].contiguous()
            self.free = self.free[:self.free.numel() - n_recv]
            tail = self.idx[self.length:self.length + n_recv]
            _lib.check(_lib.load().pysdm_b200_migration_unpack(
                recv.contiguous().data_ptr(), n_recv, slots.data_ptr(), tail.data_ptr(),
                self.multiplicity.data_ptr(), self.attributes.data_ptr(), self.n_attr,
                self.cell_origin.data_ptr(), self.position_in_cell.data_ptr(), self.n_dims,
                self.n_cap, torch.cuda.current_stream().cuda_stream), "migration_unpack")
            self._destination(tail, n_recv)  # slab-local cell ids of the arrivals
        self.length += n_recv
        self.n_received += n_recv
        return self.length


_A2A = {}


def neighbour_exchange(send, recv, rank, world, group=None):
    """the two fixed-size send/recv pairs of the neighbour path: send[0] goes to rank + 1, send[1]
    to rank - 1 (ring); recv[0] comes from rank - 1 (its send[0]), recv[1] from rank + 1 (its
    send[1]).  No counts are exchanged: each buffer's first row carries its count."""
    if world == 1:
        recv.zero_()
        return
    import torch.distributed as dist  # pylint: disable=import-outside-toplevel

    up, down = (rank + 1) % world, (rank - 1) % world
    if os.environ.get("PYSDM_MIG_EXCHANGE") == "a2a":  # one fixed-size all-to-all instead
        out = _A2A.get((send.data_ptr(), world))
        if out is None:
            out = _A2A[(send.data_ptr(), world)] = (
                torch.zeros((world,) + tuple(send.shape), dtype=send.dtype, device=send.device),
                torch.zeros((world,) + tuple(send.shape), dtype=send.dtype, device=send.device))
        buf_out, buf_in = out
        buf_out[:, :, 0, 0] = 0  # headers of the unused slots
        buf_out[up, 0].copy_(send[0])
        buf_out[down, 1].copy_(send[1])
        dist.all_to_all_single(buf_in, buf_out, group=group)
        recv[0].copy_(buf_in[down, 0])
        recv[1].copy_(buf_in[up, 1])
        return
    ops = [dist.P2POp(dist.isend, send[0], up, group), dist.P2POp(dist.irecv, recv[0], down, group),
           dist.P2POp(dist.isend, send[1], down, group), dist.P2POp(dist.irecv, recv[1], up, group)]
    for request in dist.batch_isend_irecv(ops):
        request.wait()
