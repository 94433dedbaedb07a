"""Tensor-level host bindings of the C-ABI (include/pysdm_b200.h): one Python function per
reference backend method, same names and argument order as the reference's Numba
`*Methods` bodies (PySDM/backends/impl_numba/methods/{index,pair,collisions,physics}_methods.py),
operating in place on CUDA tensors (torch is only the device-memory handle).

No CPU path: every function launches kernels of libpysdm_b200.so and raises if the
library is missing or a tensor does not live on a CUDA device.
"""

import ctypes

import torch

from . import _lib
from ._lib import PAIR_OPS, check


def _stream():
    """raw handle of torch's current stream on the current device (the private C call behind
    torch.cuda.current_stream(): ~0.3 us instead of ~5 us of Python per launch)"""
    return _raw_stream(_current_device())


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)  # pylint: disable=protected-access
_current_device = torch.cuda.current_device
if _raw_stream is None:  # (another torch: the public, slower route)
    def _stream():  # noqa: F811  pylint: disable=function-redefined
        return torch.cuda.current_stream().cuda_stream


def _p(t, dtype=None):
    if t is None:
        return None
    if not t.is_cuda:
        raise _lib.CudaLibraryError("pysdm_b200 has no CPU path: tensor is not on a CUDA device")
    assert t.is_contiguous(), "non-contiguous device view"
    if dtype is not None:
        assert t.dtype == dtype, (t.dtype, dtype)
    return t.data_ptr()


def _pi(t):
    return _p(t, torch.int64)


def _pf(t):
    return _p(t, torch.float64)


def _pb(t):
    assert t.dtype in (torch.bool, torch.uint8)
    return _p(t)


class Workspace:
    """grow-only device scratch (per key) + pinned mailbox for host results"""

    def __init__(self):
        self._dev = {}
        self._box = None

    def dev(self, key, nbytes, device):
        buf = self._dev.get((key, device))
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty(int(nbytes) + 256, dtype=torch.uint8, device=device)
            self._dev[(key, device)] = buf
        return buf

    def box(self):
        if self._box is None:
            self._box = torch.zeros(4, dtype=torch.int64).pin_memory()
        return self._box


WS = Workspace()


# ---- index methods ------------------------------------------------------------------
def identity_index(idx):
    check(_lib.load().pysdm_b200_identity_index(_pi(idx), idx.numel(), _stream()), "identity_index")


def shuffle_global(idx, length, u01):
    lib = _lib.load()
    nbytes = lib.pysdm_b200_shuffle_workspace_bytes(int(length))
    ws = WS.dev("shuffle", nbytes, idx.device)
    check(lib.pysdm_b200_shuffle_global(_pi(idx), int(length), _pf(u01), _p(ws), ws.numel(),
                                        _stream()), "shuffle_global")


_STATUS = {}


def shuffle_local(idx, u01, cell_start, length=None, max_cell=None):
    """`length` = cell_start[-1]; pass it when known on the host to avoid a device sync.
    `max_cell` = size of the largest cell, when the host knows it: populations of small cells
    then take the one-launch shared-memory shuffle instead of the general replay"""
    lib = _lib.load()
    if (max_cell is not None and max_cell <= lib.pysdm_b200_fused_max_cell_size()
            and idx.numel() < 2 ** 32 and u01.numel() >= idx.numel()):
        status = _STATUS.get(idx.device)
        if status is None:
            status = _STATUS[idx.device] = torch.zeros(1, dtype=torch.int64, device=idx.device)
        check(lib.pysdm_b200_shuffle_local_batched(
            _pi(idx), _pf(u01), _pi(cell_start), cell_start.numel() - 1, idx.numel(),
            int(max_cell), _pi(status), _stream()), "shuffle_local_batched")
        return
    if length is None:
        length = int(cell_start[-1].item())
    nbytes = lib.pysdm_b200_shuffle_workspace_bytes(int(length))
    ws = WS.dev("shuffle", nbytes, idx.device)
    check(lib.pysdm_b200_shuffle_local(_pi(idx), _pf(u01), _pi(cell_start),
                                       cell_start.numel() - 1, int(length), _p(ws), ws.numel(),
                                       _stream()), "shuffle_local")


def adaptive_presort(cell_idx, dt_left, *, cell_start, norm_factor, timestep, dv, adaptive_end,
                     skip_if_zero=None):
    """cell_idx.sort_by_key(dt_left) opening a fused adaptive sub-step, one launch
    (pysdm_b200_adaptive_presort); the sub-step that follows passes prepass_done=True.
    `norm_factor`: None when the per-cell factors are valid for the current cell_start"""
    lib = _lib.load()
    n = dt_left.numel()
    ws = WS.dev("sort_by_key", lib.pysdm_b200_sort_by_key_workspace_bytes(n), cell_idx.device)
    check(lib.pysdm_b200_adaptive_presort(
        _pi(cell_idx), _pf(dt_left), n, _p(ws), ws.numel(), _pi(skip_if_zero), _pi(cell_start),
        _pf(norm_factor), float(timestep), float(dv), _pi(adaptive_end), _stream()),
        "adaptive_presort")


SHUFFLE_STATIC_MAX_LENGTH = 262144  # PYSDM_B200_SHUFFLE_STATIC_MAX_LENGTH, include/pysdm_b200.h


def shuffle_local_static(idx, u01, cell_start, length):
    """shuffle_local through pysdm_b200_shuffle_local_static: nothing is allocated, here or in
    the library (the scratch is a fixed device array inside libpysdm_b200.so, shared by all
    calls: do not overlap them).  `length` = cell_start[-1], known on the host, at most
    SHUFFLE_STATIC_MAX_LENGTH.  Not in the reference; for callers that may not cudaMalloc"""
    if not 0 <= int(length) <= SHUFFLE_STATIC_MAX_LENGTH:
        raise ValueError(f"shuffle_local_static: length {int(length)} exceeds the static scratch "
                         f"limit of {SHUFFLE_STATIC_MAX_LENGTH} positions; use shuffle_local")
    check(_lib.load().pysdm_b200_shuffle_local_static(
        _pi(idx), _pf(u01), _pi(cell_start), cell_start.numel() - 1, int(length), _stream()),
        "shuffle_local_static")


def sort_by_key(idx, attr, skip_if_zero=None):
    """`skip_if_zero`: device int64 word; the sort does nothing when it holds 0 (speculative
    adaptive sub-steps: the working length the previous sub-step left)"""
    lib = _lib.load()
    n = attr.numel()
    ws = WS.dev("sort_by_key", lib.pysdm_b200_sort_by_key_workspace_bytes(n), idx.device)
    if skip_if_zero is None:
        check(lib.pysdm_b200_sort_by_key(_pi(idx), _pf(attr), n, _p(ws), ws.numel(), _stream()),
              "sort_by_key")
    else:
        check(lib.pysdm_b200_sort_by_key_if(_pi(idx), _pf(attr), n, _p(ws), ws.numel(),
                                            _pi(skip_if_zero), _stream()), "sort_by_key_if")


def remove_zero_n_or_flagged(multiplicity, idx, length):
    lib = _lib.load()
    ws = WS.dev("remove_zero", lib.pysdm_b200_remove_zero_workspace_bytes(int(length)), idx.device)
    box = WS.box()
    check(lib.pysdm_b200_remove_zero_n_or_flagged(
        _pi(multiplicity), _pi(idx), int(length), idx.numel(), box.data_ptr(), _p(ws), ws.numel(),
        _stream()), "remove_zero_n_or_flagged")
    return int(box[0])


def counting_sort_by_cell_id(new_idx, idx, cell_id, cell_idx, length, cell_start):
    lib = _lib.load()
    ws = WS.dev("sort_by_cell", lib.pysdm_b200_sort_by_cell_workspace_bytes(int(length)),
                idx.device)
    check(lib.pysdm_b200_sort_by_cell_id(
        _pi(new_idx), _pi(idx), _pi(cell_id), _pi(cell_idx), int(length), _pi(cell_start),
        cell_start.numel() - 1, _p(ws), ws.numel(), _stream()), "sort_by_cell_id")


def cell_id(cell_id_out, cell_origin, strides):
    n_dims, n_sd = cell_origin.shape
    strides = strides.reshape(-1).contiguous()
    check(_lib.load().pysdm_b200_cell_id(_pi(cell_id_out), _pi(cell_origin), _pi(strides), n_dims,
                                         n_sd, _stream()), "cell_id")


def invert_index(out, index):
    """out[index[c]] = c: rank -> cell map of the cell -> rank map a sort by cell used"""
    check(_lib.load().pysdm_b200_invert_index(_pi(out), _pi(index), index.numel(), _stream()),
          "invert_index")


# ---- pair methods ---------------------------------------------------------------------
def find_pairs(cell_start, is_first_in_pair, cell_id_, cell_idx, idx, length):
    check(_lib.load().pysdm_b200_find_pairs(_pi(cell_start), _pb(is_first_in_pair), _pi(cell_id_),
                                            _pi(cell_idx), _pi(idx), int(length), _stream()),
          "find_pairs")


def sort_within_pair_by_attr(idx, length, is_first_in_pair, attr):
    is_int = attr.dtype == torch.int64
    check(_lib.load().pysdm_b200_sort_within_pair_by_attr(
        _pi(idx), int(length), _pb(is_first_in_pair), _p(attr), int(is_int), _stream()),
        "sort_within_pair_by_attr")


def pair_op(op, data_out, data_in, is_first_in_pair, idx, length):
    is_int = data_in.dtype == torch.int64
    check(_lib.load().pysdm_b200_pair_op(
        PAIR_OPS[op], _pf(data_out), data_out.numel(), _p(data_in), int(is_int),
        _pb(is_first_in_pair), _pi(idx), int(length), _stream()), op + "_pair")


def sort_pair(data_out, data_in, is_first_in_pair, idx, length):
    check(_lib.load().pysdm_b200_sort_pair(_pf(data_out), data_out.numel(), _pf(data_in),
                                           _pb(is_first_in_pair), _pi(idx), int(length),
                                           _stream()), "sort_pair")


# ---- physics -----------------------------------------------------------------------------
def volume_of_water_mass(volume, mass, rho_w):
    check(_lib.load().pysdm_b200_volume_of_water_mass(_pf(volume), _pf(mass), volume.numel(),
                                                      float(rho_w), _stream()), "volume")


def mass_of_water_volume(mass, volume, rho_w):
    check(_lib.load().pysdm_b200_mass_of_water_volume(_pf(mass), _pf(volume), volume.numel(),
                                                      float(rho_w), _stream()), "mass")


# ---- collisions ----------------------------------------------------------------------------
def normalize(prob, cell_id_, cell_idx, cell_start, norm_factor, timestep, dv):
    check(_lib.load().pysdm_b200_normalize(
        _pf(prob), prob.numel(), _pi(cell_id_), _pi(cell_idx), _pi(cell_start),
        cell_start.numel() - 1, _pf(norm_factor), float(timestep), float(dv), _stream()),
        "normalize")


def compute_gamma(prob, rand, idx, length, multiplicity, cell_id_, collision_rate_deficit,
                  collision_rate, is_first_in_pair, out):
    check(_lib.load().pysdm_b200_compute_gamma(
        _pf(prob), _pf(rand), _pi(idx), int(length), _pi(multiplicity), _pi(cell_id_),
        _pi(collision_rate_deficit), _pi(collision_rate), _pb(is_first_in_pair), _pf(out),
        _stream()), "compute_gamma")


def collision_coalescence(multiplicity, idx, length, attributes, gamma, healthy, cell_id_,
                          coalescence_rate, is_first_in_pair):
    n_attr, n_sd = attributes.shape
    check(_lib.load().pysdm_b200_collision_coalescence(
        _pi(multiplicity), _pi(idx), int(length), _pf(attributes), n_attr, n_sd, _pf(gamma),
        _pi(healthy), _pi(cell_id_), _pi(coalescence_rate), _pb(is_first_in_pair), _stream()),
        "collision_coalescence")


def adaptive_sdm_end(dt_left, cell_start):
    box = WS.box()
    scratch = WS.dev("adaptive_end", 64, dt_left.device)
    check(_lib.load().pysdm_b200_adaptive_sdm_end(
        _pf(dt_left), dt_left.numel(), _pi(cell_start), box.data_ptr(), _p(scratch), _stream()),
        "adaptive_sdm_end")
    return int(box[0])


def scale_prob_for_adaptive_sdm_gamma(prob, idx, length, multiplicity, cell_id_, dt_left, dt,
                                      dt_range, is_first_in_pair, stats_n_substep, stats_dt_min):
    n_cell = dt_left.numel()
    dt_todo = WS.dev("dt_todo", 8 * n_cell, dt_left.device)
    check(_lib.load().pysdm_b200_scale_prob_for_adaptive_sdm_gamma(
        _pf(prob), _pi(idx), int(length), _pi(multiplicity), _pi(cell_id_), _pf(dt_left), n_cell,
        float(dt), float(dt_range[0]), float(dt_range[1]), _pb(is_first_in_pair),
        _pi(stats_n_substep), _pf(stats_dt_min), _p(dt_todo), _stream()),
        "scale_prob_for_adaptive_sdm_gamma")


def host_ptr(tensor):
    return ctypes.c_void_p(tensor.data_ptr())


# ---- Storage arithmetic ------------------------------------------------------------------
def elementwise(op, out, a=None, b=None, scalar=0.0):
    """out = a <op> (b or scalar); see include/pysdm_b200.h for the op table"""
    check(_lib.load().pysdm_b200_elementwise(_lib.EW[op], _pf(out), _pf(a), _pf(b), float(scalar),
                                             out.numel(), _stream()), op)


def philox_uniform(out, seed, call):
    check(_lib.load().pysdm_b200_philox_uniform(_pf(out), out.numel(), int(seed), int(call),
                                                _stream()), "philox_uniform")


def max_cell_size(cell_start):
    """largest cell_start[c+1] - cell_start[c] (host int; synchronises)"""
    box = WS.box()
    scratch = WS.dev("max_cell", 64, cell_start.device)
    check(_lib.load().pysdm_b200_max_cell_size(_pi(cell_start), cell_start.numel() - 1,
                                               box.data_ptr(), _p(scratch), _stream()),
          "max_cell_size")
    return int(box[0])


# ---- fused step ----------------------------------------------------------------------------
def fused_max_cell_size():
    return int(_lib.load().pysdm_b200_fused_max_cell_size())


def fused_collision_step(*, idx_in, idx_out, cell_start, cell_idx, cell_id_, multiplicity,
                         attributes, is_first_in_pair, gamma, norm_factor, collision_rate,
                         collision_rate_deficit, coalescence_rate, healthy, status, length,
                         max_cell, kernel, kernel_a=0.0, kernel_b=0.0, timestep, dv, rho_w,
                         substeps=1, mass_row=0, u01=None, rand=None, seed=0, call_u01=0,
                         call_rand=0, nf_slot, refresh_nf=True, gate=None, epoch=1, radius=None,
                         fall_velocity=None, berry_params=None, berry_unit=1.0, adaptive=None,
                         dependents=None, sorted_length=0, skip_coalescence=False, cache=None,
                         cache_key=None):
    """one Collision.step() in one launch; returns the CUDA error code (9 = a cell is
    larger than the fused path supports -> use the method-level functions).  kernel
    "Geometric": kernel_a = PI * collection_efficiency, `radius` and `fall_velocity` = the
    per-SD attributes of that name (f64[n_sd]); kernel "Hydrodynamic" (any `Parameterized`
    kernel): kernel_a = PI, the same two arrays, `berry_params` = the 13 parameters as passed to
    linear_collection_efficiency (ints kept as ints), `berry_unit` = its `unit`.
    `adaptive`: None, or a dict with the fields of the adaptive block of pysdm_b200_fused_args
    (dt_left, stats_n_substep, stats_dt_min, dt_min, dt_max, rank_to_cell, rank_info and
    optionally adaptive_end, length_dev): the launch is then one sub-step of the adaptive loop"""
    if (cache is not None and "args" in cache and int(max_cell) <= cache["max_cell_fused"]
            and cache.get("key") == cache_key):
        # `cache`: a dict the caller keeps for as long as every argument NOT rewritten below is
        # unchanged (same tensors, same scalars): the ~80-field argument block is built once --
        # building it costs more host time than the launch it describes
        args = cache["args"]
        args.idx_in, args.idx_out, args.cell_idx = _pi(idx_in), _pi(idx_out), _pi(cell_idx)
        args.length, args.sorted_length = int(length), int(sorted_length)
        args.max_cell_size, args.refresh_nf = int(max_cell), int(bool(refresh_nf))
        args.gate, args.epoch = _pi(gate), int(epoch)
        args.u01, args.rand = _pf(u01), _pf(rand)
        args.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        args.call_u01, args.call_rand = int(call_u01), int(call_rand)
        if adaptive is not None:
            args.adaptive_end = _pi(adaptive.get("adaptive_end"))
            args.length_dev = _pi(adaptive.get("length_dev"))
            args.prepass_done = int(bool(adaptive.get("prepass_done", False)))
        return _lib.load().pysdm_b200_fused_collision_step(ctypes.byref(args), _stream())
    n_attr, n_sd = attributes.shape
    args = _lib.FusedArgs(
        idx_in=_pi(idx_in), idx_out=_pi(idx_out), cell_start=_pi(cell_start),
        cell_idx=_pi(cell_idx), cell_id=_pi(cell_id_), multiplicity=_pi(multiplicity),
        attributes=_pf(attributes), u01=_pf(u01), rand=_pf(rand),
        is_first_in_pair=_pb(is_first_in_pair), gamma=_pf(gamma), norm_factor=_pf(norm_factor),
        collision_rate=_pi(collision_rate), collision_rate_deficit=_pi(collision_rate_deficit),
        coalescence_rate=_pi(coalescence_rate), healthy=_pi(healthy), status=_pi(status),
        length=int(length), n_sd=n_sd, n_cell=cell_start.numel() - 1, n_attr=n_attr,
        mass_row=int(mass_row), cells_per_batch=0, seed=int(seed) & 0xFFFFFFFFFFFFFFFF,
        call_u01=int(call_u01), call_rand=int(call_rand), timestep=float(timestep),
        dv=float(dv), rho_w=float(rho_w), substeps=float(substeps), kernel_a=float(kernel_a),
        kernel_b=float(kernel_b), kernel=_lib.KERNELS[kernel], reserved=0,
        max_cell_size=int(max_cell), nf_slot=_pf(nf_slot), refresh_nf=int(bool(refresh_nf)),
        gate=_pi(gate), epoch=int(epoch), radius=_pf(radius), fall_velocity=_pf(fall_velocity))
    if int(max_cell) > fused_max_cell_size():  # boxes: shuffle replay + streaming kernel(s)
        nbytes = _lib.load().pysdm_b200_fused_workspace_bytes(n_sd, cell_start.numel() - 1)
        ws = WS.dev("fused_stream", nbytes, idx_in.device)
        args.ws, args.ws_bytes = _p(ws), ws.numel()
    args.sorted_length = int(sorted_length)
    args.skip_coalescence = int(bool(skip_coalescence))
    if adaptive is not None:
        args.adaptive = 1
        args.dt_left = _pf(adaptive["dt_left"])
        args.stats_n_substep = _pi(adaptive["stats_n_substep"])
        args.stats_dt_min = _pf(adaptive["stats_dt_min"])
        args.dt_min, args.dt_max = float(adaptive["dt_min"]), float(adaptive["dt_max"])
        args.rank_to_cell = _pi(adaptive.get("rank_to_cell"))
        args.rank_info = _pi(adaptive.get("rank_info"))
        args.adaptive_end = _pi(adaptive.get("adaptive_end"))
        args.length_dev = _pi(adaptive.get("length_dev"))
        args.prepass_done = int(bool(adaptive.get("prepass_done", False)))
    if dependents is not None:  # see pysdm_b200_fused_args `dep_*`
        args.dep_volume = _pf(dependents.get("volume"))
        args.dep_radius = _pf(dependents.get("radius"))
        args.dep_velocity = _pf(dependents.get("velocity"))
        args.dep_table_b = _pf(dependents.get("table_b"))
        args.dep_table_c = _pf(dependents.get("table_c"))
        args.dep_radius_scale = float(dependents.get("radius_scale", 0.0))
        args.dep_radius_exponent = float(dependents.get("radius_exponent", 0.0))
        args.dep_table_factor = float(dependents.get("table_factor", 0.0))
        args.dep_radius_max = float(dependents.get("radius_max", float("inf")))
        args.dep_status = _pi(dependents.get("status"))
    if berry_params is not None:
        assert len(berry_params) == 13
        args.berry_params = (ctypes.c_double * 13)(*[float(v) for v in berry_params])
        args.berry_int_mask = int_mask_of(berry_params)
        args.berry_unit = float(berry_unit)
    if cache is not None:  # (the tensors behind the cached pointers are the caller's to keep)
        cache["args"], cache["max_cell_fused"] = args, fused_max_cell_size()
        cache["key"] = cache_key  # (what the caller says identifies everything not rewritten)
        if int(max_cell) > cache["max_cell_fused"]:
            cache.pop("args")  # the streaming step's workspace may move: no caching
    return _lib.load().pysdm_b200_fused_collision_step(ctypes.byref(args), _stream())


def fused_breakup(*, multiplicity, idx, attributes, gamma, is_first_in_pair, cell_id_, healthy,
                  coalescence_rate, breakup_rate, breakup_rate_deficit, length, mass_row,
                  ec_kind, nf_kind, ec_const=0.0, eb_const=1.0, nf_n=1.0, ec_params=None,
                  ec_unit=1.0, vmin=0.0, nfmax=None, handle_all_breakups=False, max_multiplicity,
                  constants, rand_proc=None, rand_frag=None, seed=0, call_proc=0, call_frag=0,
                  volume=None, radius=None, fall_velocity=None, dependents=None,
                  count_overflows=False):
    """the breakup half of a step in one launch (pysdm_b200_fused_breakup); `constants`: dict with
    rho_w, PI, sgm_w, uJ as the Formulae / `si` carry them.  Returns the number of overflowing
    pairs when `count_overflows` (synchronises), else None"""
    n_attr, n_sd = attributes.shape
    counter = None
    if count_overflows:
        counter = torch.zeros(1, dtype=torch.int64, device=multiplicity.device)
    pi, rho_w, sgm_w = constants["PI"], constants["rho_w"], constants["sgm_w"]
    args = _lib.BreakupArgs(
        multiplicity=_pi(multiplicity), idx=_pi(idx), attributes=_pf(attributes), gamma=_pf(gamma),
        is_first_in_pair=_pb(is_first_in_pair), cell_id=_pi(cell_id_), healthy=_pi(healthy),
        coalescence_rate=_pi(coalescence_rate), breakup_rate=_pi(breakup_rate),
        breakup_rate_deficit=_pi(breakup_rate_deficit), n_overflow=_pi(counter),
        length=int(length), n_sd=n_sd, n_attr=n_attr, mass_row=int(mass_row),
        rand_proc=_pf(rand_proc), rand_frag=_pf(rand_frag),
        seed=int(seed) & 0xFFFFFFFFFFFFFFFF, call_proc=int(call_proc), call_frag=int(call_frag),
        volume=_pf(volume), radius=_pf(radius), fall_velocity=_pf(fall_velocity),
        ec_kind=int(ec_kind), nf_kind=int(nf_kind), ec_const=float(ec_const),
        eb_const=float(eb_const), nf_n=float(nf_n), ec_unit=float(ec_unit), vmin=float(vmin),
        nfmax=float(0.0 if nfmax is None else nfmax), has_nfmax=int(nfmax is not None),
        handle_all_breakups=int(bool(handle_all_breakups)),
        max_multiplicity=float(max_multiplicity),
        # products formed exactly as the reference's Python forms them
        rho_w=float(rho_w), six_over_pi=6 / pi, pi_sgm_w=pi * sgm_w, two_thirds=2 / 3,
        nf_sc_factor=pi * sgm_w * (6 / pi) ** (2 / 3), half_rho_w=rho_w / 2,
        micro_joule=float(constants["uJ"]))
    if ec_params is not None:
        assert len(ec_params) == 13
        args.ec_params = (ctypes.c_double * 13)(*[float(v) for v in ec_params])
        args.ec_int_mask = int_mask_of(ec_params)
    if dependents is not None:
        args.dep_volume = _pf(dependents.get("volume"))
        args.dep_radius = _pf(dependents["radius"])
        args.dep_velocity = _pf(dependents.get("velocity"))
        args.dep_table_b = _pf(dependents.get("table_b"))
        args.dep_table_c = _pf(dependents.get("table_c"))
        args.dep_radius_scale = float(dependents["radius_scale"])
        args.dep_radius_exponent = float(dependents["radius_exponent"])
        args.dep_table_factor = float(dependents.get("table_factor", 0.0))
        args.dep_radius_max = float(dependents.get("radius_max", float("inf")))
        args.dep_status = _pi(dependents.get("status"))
    check(_lib.load().pysdm_b200_fused_breakup(ctypes.byref(args), _stream()), "fused_breakup")
    return int(counter.item()) if count_overflows else None


def dependent_attributes(mass, rho_w, *, radius_scale, radius_exponent, table_factor=0.0,
                         radius_max=float("inf"), table_b=None, table_c=None, volume=None,
                         radius=None, velocity=None, status=None):
    """volume / radius / Gunn-Kinzer fall velocity of every super-droplet in one pass"""
    check(_lib.load().pysdm_b200_dependent_attributes(
        _pf(mass), mass.numel(), float(rho_w), float(radius_scale), float(radius_exponent),
        float(table_factor), float(radius_max), _pf(table_b), _pf(table_c), _pf(volume),
        _pf(radius), _pf(velocity), _pi(status), _stream()), "dependent_attributes")


# ---- terminal velocities, collection efficiency (csrc/physics.cu) ---------------------------
def _host_f64(values):
    """small host-side parameter arrays (kept alive by the caller's frame during the call)"""
    arr = (ctypes.c_double * len(values))(*[float(v) for v in values])
    return arr


def interpolation(output, radius, factor, b, c):
    check(_lib.load().pysdm_b200_interpolation(_pf(output), _pf(radius), radius.numel(),
                                               float(factor), _pf(b), _pf(c), _stream()),
          "interpolation")


def terminal_velocity_rogers_yau(values, radius, params):
    """params = (small_k, medium_k, large_k, small_r_limit, medium_r_limit)"""
    check(_lib.load().pysdm_b200_terminal_velocity_rogers_yau(
        _pf(values), _pf(radius), values.numel(), *[float(p) for p in params], _stream()),
        "terminal_velocity")


def power_series(values, radius, prefactors, powers):
    pre, pw = _host_f64(prefactors), _host_f64(powers)
    check(_lib.load().pysdm_b200_power_series(_pf(values), _pf(radius), values.numel(), len(pw),
                                              pre, pw, _stream()), "power_series")


def int_mask_of(params):
    """bit e set when params[e] is an integer (include/pysdm_b200.h)"""
    import numbers  # pylint: disable=import-outside-toplevel

    return sum(1 << e for e, v in enumerate(params)
               if isinstance(v, numbers.Integral) and not isinstance(v, bool))


def linear_collection_efficiency(params, int_mask, output, radii, is_first_in_pair, idx, length,
                                 unit):
    par = _host_f64(params)
    assert len(par) == 13
    check(_lib.load().pysdm_b200_linear_collection_efficiency(
        par, int(int_mask), _pf(output), output.numel(), _pf(radii), _pb(is_first_in_pair),
        _pi(idx), int(length), float(unit), _stream()), "linear_collection_efficiency")


# ---- fragmentation functions (csrc/physics.cu) ---------------------------------------------
def _lim(vmin, nfmax):
    return float(vmin), int(nfmax is not None), float(0.0 if nfmax is None else nfmax)


def slams_fragmentation(n_fragment, frag_volume, x_plus_y, probs, rand, vmin, nfmax):
    check(_lib.load().pysdm_b200_slams_fragmentation(
        _pf(n_fragment), _pf(frag_volume), _pf(x_plus_y), _pf(probs), _pf(rand), rand.numel(),
        *_lim(vmin, nfmax), _stream()), "slams_fragmentation")


def exp_fragmentation(n_fragment, scale, frag_volume, x_plus_y, rand, vmin, nfmax, tol=1e-5):
    check(_lib.load().pysdm_b200_exp_fragmentation(
        _pf(n_fragment), float(scale), _pf(frag_volume), _pf(x_plus_y), _pf(rand), rand.numel(),
        *_lim(vmin, nfmax), float(tol), _stream()), "exp_fragmentation")


def feingold1988_fragmentation(n_fragment, scale, frag_volume, x_plus_y, rand, fragtol, vmin,
                               nfmax):
    check(_lib.load().pysdm_b200_feingold1988_fragmentation(
        _pf(n_fragment), float(scale), _pf(frag_volume), _pf(x_plus_y), _pf(rand), rand.numel(),
        float(fragtol), *_lim(vmin, nfmax), _stream()), "feingold1988_fragmentation")


def gauss_fragmentation(n_fragment, mu, sigma, frag_volume, x_plus_y, rand, vmin, nfmax):
    check(_lib.load().pysdm_b200_gauss_fragmentation(
        _pf(n_fragment), float(mu), float(sigma), _pf(frag_volume), _pf(x_plus_y), _pf(rand),
        rand.numel(), *_lim(vmin, nfmax), _stream()), "gauss_fragmentation")


def straub_fragmentation(n_fragment, CW, gam, ds, frag_volume, v_max, x_plus_y, rand, vmin,
                         nfmax, Nr1, Nr2, Nr3, Nr4, Nrt, d34):
    check(_lib.load().pysdm_b200_straub_fragmentation(
        _pf(n_fragment), _pf(CW), _pf(gam), _pf(ds), _pf(frag_volume), _pf(v_max), _pf(x_plus_y),
        _pf(rand), _pf(Nr1), _pf(Nr2), _pf(Nr3), _pf(Nr4), _pf(Nrt), _pf(d34), rand.numel(),
        *_lim(vmin, nfmax), _stream()), "straub_fragmentation")


def ll82_fragmentation(n_fragment, CKE, W, W2, St, ds, dl, dcoal, frag_volume, x_plus_y, rand,
                       vmin, nfmax, Rf, Rs, Rd, tol=1e-8):
    check(_lib.load().pysdm_b200_ll82_fragmentation(
        _pf(n_fragment), _pf(CKE), _pf(W), _pf(W2), _pf(St), _pf(ds), _pf(dl), _pf(dcoal),
        _pf(frag_volume), _pf(x_plus_y), _pf(rand), _pf(Rf), _pf(Rs), _pf(Rd), rand.numel(),
        float(tol), *_lim(vmin, nfmax), _stream()), "ll82_fragmentation")


def ll82_coalescence_check(Ec, dl):
    check(_lib.load().pysdm_b200_ll82_coalescence_check(_pf(Ec), _pf(dl), Ec.numel(), _stream()),
          "ll82_coalescence_check")


# ---- breakup (csrc/breakup.cu) -----------------------------------------------------------------
def collision_coalescence_breakup(multiplicity, idx, length, attributes, gamma, rand, Ec, Eb,
                                  fragment_mass, healthy, cell_id_, coalescence_rate,
                                  breakup_rate, breakup_rate_deficit, is_first_in_pair,
                                  max_multiplicity, particle_mass, handle_all_breakups=False,
                                  count_overflows=False):
    """returns the number of pairs that raised the overflow flag when `count_overflows`
    (synchronises), else None"""
    n_attr, n_sd = attributes.shape
    counter = None
    if count_overflows:
        counter = torch.zeros(1, dtype=torch.int64, device=multiplicity.device)
    check(_lib.load().pysdm_b200_collision_coalescence_breakup(
        _pi(multiplicity), _pi(idx), int(length), _pf(attributes), n_attr, n_sd, _pf(gamma),
        _pf(rand), _pf(Ec), _pf(Eb), _pf(fragment_mass), _pi(healthy), _pi(cell_id_),
        _pi(coalescence_rate), _pi(breakup_rate), _pi(breakup_rate_deficit),
        _pb(is_first_in_pair), float(max_multiplicity), _pf(particle_mass),
        int(bool(handle_all_breakups)), _pi(counter), _stream()),
        "collision_coalescence_breakup")
    return int(counter.item()) if count_overflows else None


# ---- products and transport side (csrc/transport.cu) ------------------------------
def _is_int(x):
    import numbers  # pylint: disable=import-outside-toplevel

    return int(isinstance(x, numbers.Integral) and not isinstance(x, bool))


def moments(moment_0, moments_, multiplicity, attr_data, cell_id_, idx, length, ranks, min_x,
            max_x, x_attr, weighting_attribute, weighting_rank, skip_division_by_m0):
    """`ranks`: host sequence of floats (the reference's `ranks` storage holds floats)"""
    ranks = [float(r) for r in ranks]
    check(_lib.load().pysdm_b200_moments(
        _pf(moment_0), _pf(moments_), _pi(multiplicity), _pf(attr_data) if ranks else None,
        _pi(cell_id_), _pi(idx), int(length), _host_f64(ranks) if ranks else None, len(ranks),
        float(min_x), float(max_x), _pf(x_attr), _pf(weighting_attribute), float(weighting_rank),
        _is_int(weighting_rank), int(bool(skip_division_by_m0)), moment_0.numel(), _stream()),
        "moments")


def spectrum_moments(moment_0, moments_, multiplicity, attr_data, cell_id_, idx, length, rank,
                     x_bins, x_attr, weighting_attribute, weighting_rank):
    n_bins, n_cell = moments_.shape
    assert x_bins.numel() == n_bins + 1 and tuple(moment_0.shape) == (n_bins, n_cell)
    check(_lib.load().pysdm_b200_spectrum_moments(
        _pf(moment_0), _pf(moments_), _pi(multiplicity), _pf(attr_data), _pi(cell_id_), _pi(idx),
        int(length), float(rank), _is_int(rank), _pf(x_bins), n_bins, _pf(x_attr),
        _pf(weighting_attribute), float(weighting_rank), _is_int(weighting_rank), n_cell,
        _stream()), "spectrum_moments")


def calculate_displacement(dim, displacement, courant, cell_origin, position_in_cell, n_substeps,
                           implicit_in_space=True):
    n_dims, n_sd = displacement.shape
    assert courant.dim() == n_dims
    shape = (ctypes.c_int64 * n_dims)(*courant.shape)
    check(_lib.load().pysdm_b200_calculate_displacement(
        int(dim), n_dims, int(bool(implicit_in_space)), _pf(displacement), _pf(courant), shape,
        _pi(cell_origin), _pf(position_in_cell), n_sd, float(n_substeps), _stream()),
        "calculate_displacement")


def displacement_step(displacement, courant, cell_origin, position_in_cell, grid, n_substeps=1,
                      implicit_in_space=True):
    """Displacement.__call__ without sedimentation in one launch (pysdm_b200_displacement_step):
    `courant` = the components of the Courant field (device tensors), `grid` = host ints"""
    n_dims, n_sd = displacement.shape
    assert len(courant) == n_dims == len(grid)
    pointers = (ctypes.c_void_p * n_dims)(*[_pf(c) for c in courant])
    shapes = (ctypes.c_int64 * (n_dims * n_dims))(*[int(v) for c in courant for v in c.shape])
    check(_lib.load().pysdm_b200_displacement_step(
        n_dims, int(bool(implicit_in_space)), _pf(displacement), pointers, shapes,
        (ctypes.c_int64 * n_dims)(*[int(g) for g in grid]), _pi(cell_origin),
        _pf(position_in_cell), n_sd, float(n_substeps), _stream()), "displacement_step")


def flag_precipitated(cell_origin, position_in_cell, water_mass, multiplicity, idx, length,
                      healthy, precipitation_counting_level_index, displacement):
    """returns the rainfall mass (host float; synchronises)"""
    rain = torch.zeros(1, dtype=torch.float64, device=idx.device)
    check(_lib.load().pysdm_b200_flag_precipitated(
        _pi(cell_origin[-1]), _pf(position_in_cell[-1]), _pf(water_mass), _pi(multiplicity),
        _pi(idx), int(length), idx.numel(), _pi(healthy),
        float(precipitation_counting_level_index), _pf(displacement[-1]), _pf(rain), _stream()),
        "flag_precipitated")
    return float(rain.item())


def flag_out_of_column(cell_origin, position_in_cell, idx, length, healthy,
                       domain_top_level_index):
    check(_lib.load().pysdm_b200_flag_out_of_column(
        _pi(cell_origin[-1]), _pf(position_in_cell[-1]), _pi(idx), int(length), idx.numel(),
        _pi(healthy), float(domain_top_level_index), _stream()), "flag_out_of_column")


def floor_to_int(out, src):
    check(_lib.load().pysdm_b200_floor_to_i64(_pi(out), _pf(src), src.numel(), _stream()), "floor")


def sub_int(inout, b):
    check(_lib.load().pysdm_b200_sub_i64(_pf(inout), _pi(b), b.numel(), _stream()), "sub_int")


def row_modulo_int(out, divisor):
    n_rows, n_cols = out.shape if out.dim() == 2 else (1, out.numel())
    check(_lib.load().pysdm_b200_row_modulo_i64(_pi(out), _pi(divisor), n_rows, n_cols,
                                                _stream()), "row_modulo")


def seeding(idx, multiplicity, extensive_attributes, seeded_particle_index,
            seeded_particle_multiplicity, seeded_particle_extensive_attributes,
            number_of_super_particles_to_inject):
    """seeding_methods.py:12-65; the reference's closing assert (as many injected as asked
    for) is raised here from the count the kernel reports"""
    lib = _lib.load()
    n_attr, n_sd = extensive_attributes.shape
    n_seeds = seeded_particle_multiplicity.numel()
    assert seeded_particle_extensive_attributes.shape == (n_attr, n_seeds)
    ws = WS.dev("seeding", lib.pysdm_b200_seeding_workspace_bytes(int(n_sd)), idx.device)
    box = WS.box()
    check(lib.pysdm_b200_seeding(
        _pi(idx), _pi(multiplicity), _pf(extensive_attributes), int(n_attr), int(n_sd),
        _pi(seeded_particle_index), _pi(seeded_particle_multiplicity),
        _pf(seeded_particle_extensive_attributes), int(n_seeds),
        int(number_of_super_particles_to_inject), box.data_ptr(), _p(ws), ws.numel(), _stream()),
        "seeding")
    assert int(box[0]) == int(number_of_super_particles_to_inject)
