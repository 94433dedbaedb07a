"""GPU: `pysdm_b200_shuffle_local_static` (include/pysdm_b200.h) -- shuffle_local for callers that
may not allocate: no workspace argument, the scratch is a fixed device array inside the library.
Called through the C-ABI and checked, bit for bit (index work), against the existing entry point
`pysdm_b200_shuffle_local` on the same inputs, on several calls in a row (the array keeps the
previous call's contents), and against the oracle's serial swap chain on a small case.

The file sorts last on purpose (pytest collects alphabetically, the round's end runs `-x`).
"""

import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_LENGTH = 262144  # PYSDM_B200_SHUFFLE_STATIC_MAX_LENGTH
CUDA_ERROR_INVALID_VALUE = 1


def _population(torch, seed, n_sd, n_cell, ragged):
    """idx, u01, cell_start of a population sorted by cell (host construction, then upload)"""
    rng = np.random.default_rng(seed)
    if ragged:
        sizes = rng.multinomial(n_sd, rng.dirichlet(np.full(n_cell, 0.3)))  # empty cells too
    else:
        sizes = np.full(n_cell, n_sd // n_cell)
        sizes[: n_sd - sizes.sum()] += 1
    cell_start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    assert cell_start[-1] == n_sd
    idx = rng.permutation(n_sd).astype(np.int64)
    u01 = rng.random(n_sd)
    dev = torch.device("cuda:0")
    return (torch.from_numpy(idx).to(dev), torch.from_numpy(u01).to(dev),
            torch.from_numpy(cell_start).to(dev), idx, u01, cell_start)


def _with_workspace(torch, lib, idx, u01, cell_start, length):
    """the existing entry point, caller-provided workspace"""
    out = idx.clone()
    nbytes = lib.pysdm_b200_shuffle_workspace_bytes(length)
    ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=idx.device)
    err = lib.pysdm_b200_shuffle_local(out.data_ptr(), u01.data_ptr(), cell_start.data_ptr(),
                                       cell_start.numel() - 1, length, ws.data_ptr(), ws.numel(),
                                       None)
    assert err == 0, err
    torch.cuda.synchronize()
    return out


def _static(torch, lib, idx, u01, cell_start, length):
    """the new entry point; checks that torch's allocator and the driver saw no allocation"""
    out = idx.clone()
    torch.cuda.synchronize()
    reserved, allocated = torch.cuda.memory_reserved(), torch.cuda.memory_allocated()
    free_before, _ = torch.cuda.mem_get_info()
    err = lib.pysdm_b200_shuffle_local_static(out.data_ptr(), u01.data_ptr(),
                                              cell_start.data_ptr(), cell_start.numel() - 1,
                                              length, None)
    assert err == 0, err
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_reserved() == reserved and torch.cuda.memory_allocated() == allocated
    assert free_after >= free_before, (free_before, free_after)
    return out


# (seed, n_sd, n_cell, ragged): a large first call, then smaller and differently shaped ones that
# find the array full of the previous call's lists and counters, then the maximum size
CALLS = [(1, 200_000, 1500, True), (2, 4099, 7, True), (3, 65_536, 1, False),
         (4, 30_000, 30_000, False), (5, MAX_LENGTH, 2048, False), (6, 1000, 3, True)]


def test_static_scratch_equals_workspace_entry_point():
    import torch  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import _lib  # pylint: disable=import-outside-toplevel

    lib = _lib.load()
    # set-up call: length 0 only looks the array's device address up
    assert lib.pysdm_b200_shuffle_local_static(None, None, None, 0, 0, None) == 0
    for seed, n_sd, n_cell, ragged in CALLS:
        idx, u01, cell_start, *_ = _population(torch, seed, n_sd, n_cell, ragged)
        expected = _with_workspace(torch, lib, idx, u01, cell_start, n_sd)
        got = _static(torch, lib, idx, u01, cell_start, n_sd)
        assert torch.equal(got, expected), (seed, n_sd, n_cell)
        assert torch.equal(torch.sort(got).values, torch.arange(n_sd, device=got.device))
        # the same call again, straight after: same answer from a dirty array
        again = _static(torch, lib, idx, u01, cell_start, n_sd)
        assert torch.equal(again, expected), (seed, n_sd, n_cell, "second call")


def test_static_scratch_equals_oracle_swap_chain():
    import torch  # pylint: disable=import-outside-toplevel

    from oracle import sdm_oracle  # pylint: disable=import-outside-toplevel
    from pysdm_b200 import _lib  # pylint: disable=import-outside-toplevel

    lib = _lib.load()
    for seed, n_sd, n_cell in ((11, 5000, 40), (12, 777, 5)):
        idx, u01, cell_start, idx_h, u01_h, cell_start_h = _population(torch, seed, n_sd, n_cell,
                                                                        True)
        expected = idx_h.copy()
        sdm_oracle.shuffle_local(expected, u01_h, cell_start_h)
        got = _static(torch, lib, idx, u01, cell_start, n_sd)
        np.testing.assert_array_equal(got.cpu().numpy(), expected)


def test_static_scratch_rejects_lengths_beyond_the_maximum(capfd):
    import torch  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import _lib, methods  # pylint: disable=import-outside-toplevel

    lib = _lib.load()
    assert methods.SHUFFLE_STATIC_MAX_LENGTH == MAX_LENGTH
    n_sd = MAX_LENGTH + 1
    idx, u01, cell_start, *_ = _population(torch, 21, n_sd, 16, False)
    before = idx.clone()
    err = lib.pysdm_b200_shuffle_local_static(idx.data_ptr(), u01.data_ptr(),
                                              cell_start.data_ptr(), 16, n_sd, None)
    torch.cuda.synchronize()
    assert err == CUDA_ERROR_INVALID_VALUE
    assert "pysdm_b200_shuffle_local_static" in capfd.readouterr().err
    assert torch.equal(idx, before)  # nothing ran
    assert lib.pysdm_b200_shuffle_local_static(idx.data_ptr(), u01.data_ptr(),
                                               cell_start.data_ptr(), 16, -1, None) \
        == CUDA_ERROR_INVALID_VALUE
    with pytest.raises(ValueError, match="static scratch limit"):
        methods.shuffle_local_static(idx, u01, cell_start, n_sd)
    # the library is still usable afterwards, through the host-side mirror as well
    idx, u01, cell_start, *_ = _population(torch, 22, 10_000, 100, True)
    expected = _with_workspace(torch, lib, idx, u01, cell_start, 10_000)
    methods.shuffle_local_static(idx, u01, cell_start, 10_000)
    torch.cuda.synchronize()
    assert torch.equal(idx, expected)
    assert ctypes.sizeof(ctypes.c_void_p) == 8
