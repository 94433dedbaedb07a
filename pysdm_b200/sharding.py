"""Cell sharding for multi-GPU runs: collisions only pair super-droplets of
one cell, so contiguous ranges of cells go to ranks and the collision step needs no
communication.  Host-side logic only (pure Python/numpy): which cells and which
super-droplets a rank owns, and how per-cell results are put back together.
`torch.distributed` appears only in `gather_per_cell` and `reduce_products` (products /
diagnostics)."""

import numpy as np


def cell_range(n_cell, rank, world_size):
    """contiguous, balanced [begin, end) range of cell ids owned by `rank`"""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    base, extra = divmod(int(n_cell), int(world_size))
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def owner_of_cell(cell_id, n_cell, world_size):
    """rank owning each cell id (vectorised inverse of `cell_range`)"""
    cell_id = np.asarray(cell_id, dtype=np.int64)
    base, extra = divmod(int(n_cell), int(world_size))
    split = extra * (base + 1)  # cells below `split` live on the `extra` larger shards
    big = cell_id // max(base + 1, 1)
    small = extra + (cell_id - split) // max(base, 1)
    return np.where(cell_id < split, big, small).astype(np.int64)


def shard_population(multiplicity, attributes, cell_id, n_cell, rank, world_size):
    """the super-droplets of this rank's cells, cell ids renumbered to start at 0;
    returns (multiplicity, attributes[A, n_local], local_cell_id, n_local_cell, global_index)"""
    begin, end = cell_range(n_cell, rank, world_size)
    cell_id = np.asarray(cell_id, dtype=np.int64)
    mine = np.flatnonzero((cell_id >= begin) & (cell_id < end))
    attributes = np.atleast_2d(np.asarray(attributes))
    return (np.asarray(multiplicity)[mine], attributes[:, mine], cell_id[mine] - begin,
            end - begin, mine)


def gather_per_cell(local_values, n_cell, group=None):
    """all-gather of a per-cell array (counters, rates): every rank gets the global array.
    `local_values`: 1-D torch tensor of this rank's cells (CPU with gloo, CUDA with NCCL)."""
    import torch  # pylint: disable=import-outside-toplevel
    import torch.distributed as dist  # pylint: disable=import-outside-toplevel

    world = dist.get_world_size(group)
    sizes = [cell_range(n_cell, r, world)[1] - cell_range(n_cell, r, world)[0]
             for r in range(world)]
    pad = max(sizes)
    buf = torch.zeros(pad, dtype=local_values.dtype, device=local_values.device)
    buf[:local_values.numel()] = local_values
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    return torch.cat([o[:n] for o, n in zip(out, sizes)])


def reduce_products(sums=None, minima=None, maxima=None, group=None):
    """domain-wide products of a sharded run (SURVEY.md 8e ii): every rank passes the values of
    its own cells and gets the global ones -- `sums` (rate totals, moments over the domain,
    super-droplet counts), `minima` (`stats_dt_min`, `healthy`: one unhealthy rank makes the
    run unhealthy), `maxima` (`stats_n_substep`).  Each argument: a torch tensor or None; one
    all-reduce per argument (CPU tensors with gloo, CUDA tensors with NCCL), in place.
    Returns the tuple (sums, minima, maxima)."""
    import torch.distributed as dist  # pylint: disable=import-outside-toplevel

    if not dist.is_initialized():  # a single-process run: its own values are the global ones
        return sums, minima, maxima
    for tensor, op in ((sums, dist.ReduceOp.SUM), (minima, dist.ReduceOp.MIN),
                       (maxima, dist.ReduceOp.MAX)):
        if tensor is not None and tensor.numel():
            dist.all_reduce(tensor, op=op, group=group)
    return sums, minima, maxima
