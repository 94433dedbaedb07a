"""ctypes binding of libpysdm_b200.so (the C-ABI declared in include/pysdm_b200.h).

There is NO fallback: if the CUDA library is missing or no CUDA device is present the
backend cannot be constructed and says so.
"""

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# PYSDM_B200_SO: tuning variants of the same sources (see pysdm_b200/build.py --out/-D)
SO_PATH = os.environ.get("PYSDM_B200_SO") or os.path.join(_HERE, "libpysdm_b200.so")

P = ctypes.c_void_p
I64 = ctypes.c_int64
U64 = ctypes.c_uint64
U32 = ctypes.c_uint32
F64 = ctypes.c_double
INT = ctypes.c_int
SIZE = ctypes.c_size_t

# name -> (restype, argtypes); mirrors include/pysdm_b200.h one to one
SIGNATURES = {
    "pysdm_b200_identity_index": (INT, (P, I64, P)),
    "pysdm_b200_shuffle_workspace_bytes": (SIZE, (I64,)),
    "pysdm_b200_shuffle_global": (INT, (P, I64, P, P, SIZE, P)),
    "pysdm_b200_shuffle_local": (INT, (P, P, P, I64, I64, P, SIZE, P)),
    "pysdm_b200_shuffle_local_static": (INT, (P, P, P, I64, I64, P)),
    "pysdm_b200_sort_by_cell_workspace_bytes": (SIZE, (I64,)),
    "pysdm_b200_sort_by_cell_id": (INT, (P, P, P, P, I64, P, I64, P, SIZE, P)),
    "pysdm_b200_sort_by_key_workspace_bytes": (SIZE, (I64,)),
    "pysdm_b200_sort_by_key": (INT, (P, P, I64, P, SIZE, P)),
    "pysdm_b200_sort_by_key_if": (INT, (P, P, I64, P, SIZE, P, P)),
    "pysdm_b200_adaptive_presort": (INT, (P, P, I64, P, SIZE, P, P, P, F64, F64, P, P)),
    "pysdm_b200_remove_zero_workspace_bytes": (SIZE, (I64,)),
    "pysdm_b200_remove_zero_n_or_flagged": (INT, (P, P, I64, I64, P, P, SIZE, P)),
    "pysdm_b200_cell_id": (INT, (P, P, P, I64, I64, P)),
    "pysdm_b200_find_pairs": (INT, (P, P, P, P, P, I64, P)),
    "pysdm_b200_sort_within_pair_by_attr": (INT, (P, I64, P, P, INT, P)),
    "pysdm_b200_pair_op": (INT, (INT, P, I64, P, INT, P, P, I64, P)),
    "pysdm_b200_sort_pair": (INT, (P, I64, P, P, P, I64, P)),
    "pysdm_b200_normalize": (INT, (P, I64, P, P, P, I64, P, F64, F64, P)),
    "pysdm_b200_compute_gamma": (INT, (P, P, P, I64, P, P, P, P, P, P, P)),
    "pysdm_b200_scale_prob_for_adaptive_sdm_gamma": (
        INT, (P, P, I64, P, P, P, I64, F64, F64, F64, P, P, P, P, P)),
    "pysdm_b200_adaptive_sdm_end": (INT, (P, I64, P, P, P, P)),
    "pysdm_b200_collision_coalescence": (INT, (P, P, I64, P, I64, I64, P, P, P, P, P, P)),
    "pysdm_b200_elementwise": (INT, (INT, P, P, P, F64, I64, P)),
    "pysdm_b200_row_modulo": (INT, (P, P, I64, I64, P)),
    "pysdm_b200_reduce_scratch_bytes": (SIZE, ()),
    "pysdm_b200_minmax_f64": (INT, (P, I64, INT, P, P, P)),
    "pysdm_b200_minmax_i64": (INT, (P, I64, INT, P, P, P)),
    "pysdm_b200_philox_uniform": (INT, (P, I64, U64, U64, P)),
    "pysdm_b200_philox_raw": (INT, (P, U32, U32, U32, U32, U32, U32, P)),
    "pysdm_b200_volume_of_water_mass": (INT, (P, P, I64, F64, P)),
    "pysdm_b200_mass_of_water_volume": (INT, (P, P, I64, F64, P)),
    "pysdm_b200_max_cell_size": (INT, (P, I64, P, P, P)),
    "pysdm_b200_fused_max_cell_size": (I64, ()),
    "pysdm_b200_fused_workspace_bytes": (SIZE, (I64, I64)),
    "pysdm_b200_fused_collision_step": (INT, (P, P)),
    "pysdm_b200_fused_breakup": (INT, (P, P)),
    "pysdm_b200_shuffle_local_batched": (INT, (P, P, P, I64, I64, I64, P, P)),
    "pysdm_b200_invert_index": (INT, (P, P, I64, P)),
    "pysdm_b200_dependent_attributes": (INT, (P, I64, F64, F64, F64, F64, F64, P, P, P, P, P, P, P)),
    "pysdm_b200_interpolation": (INT, (P, P, I64, F64, P, P, P)),
    "pysdm_b200_terminal_velocity_rogers_yau": (INT, (P, P, I64, F64, F64, F64, F64, F64, P)),
    "pysdm_b200_power_series": (INT, (P, P, I64, I64, P, P, P)),
    "pysdm_b200_linear_collection_efficiency": (INT, (P, I64, P, I64, P, P, P, I64, F64, P)),
    "pysdm_b200_slams_fragmentation": (INT, (P, P, P, P, P, I64, F64, INT, F64, P)),
    "pysdm_b200_exp_fragmentation": (INT, (P, F64, P, P, P, I64, F64, INT, F64, F64, P)),
    "pysdm_b200_feingold1988_fragmentation": (
        INT, (P, F64, P, P, P, I64, F64, F64, INT, F64, P)),
    "pysdm_b200_gauss_fragmentation": (INT, (P, F64, F64, P, P, P, I64, F64, INT, F64, P)),
    "pysdm_b200_straub_fragmentation": (
        INT, (P,) * 14 + (I64, F64, INT, F64, P)),
    "pysdm_b200_ll82_fragmentation": (INT, (P,) * 14 + (I64, F64, F64, INT, F64, P)),
    "pysdm_b200_ll82_coalescence_check": (INT, (P, P, I64, P)),
    "pysdm_b200_floor_to_i64": (INT, (P, P, I64, P)),
    "pysdm_b200_sub_i64": (INT, (P, P, I64, P)),
    "pysdm_b200_row_modulo_i64": (INT, (P, P, I64, I64, P)),
    "pysdm_b200_moments": (
        INT, (P, P, P, P, P, P, I64, P, I64, F64, F64, P, P, F64, INT, INT, I64, P)),
    "pysdm_b200_spectrum_moments": (
        INT, (P, P, P, P, P, P, I64, F64, INT, P, I64, P, P, F64, INT, I64, P)),
    "pysdm_b200_calculate_displacement": (INT, (INT, INT, INT, P, P, P, P, P, I64, F64, P)),
    "pysdm_b200_displacement_step": (INT, (ctypes.c_int, ctypes.c_int, P, P, P, P, P, P, I64, F64, P)),
    "pysdm_b200_flag_precipitated": (INT, (P, P, P, P, P, I64, I64, P, F64, P, P, P)),
    "pysdm_b200_flag_out_of_column": (INT, (P, P, P, I64, I64, P, F64, P)),
    "pysdm_b200_seeding_workspace_bytes": (SIZE, (I64,)),
    "pysdm_b200_seeding": (INT, (P, P, P, I64, I64, P, P, P, I64, I64, P, P, SIZE, P)),
    "pysdm_b200_migration_destination": (INT, (P, P, P, P, I64, I64, INT, P, I64, P)),
    "pysdm_b200_migration_pack": (INT, (P, I64, P, I64, I64, P, P, I64, P, P, INT, I64, P)),
    "pysdm_b200_migration_unpack": (INT, (P, I64, P, P, P, P, I64, P, P, INT, I64, P)),
    "pysdm_b200_migration_split_workspace_bytes": (SIZE, (I64,)),
    "pysdm_b200_migration_split": (
        INT, (P, I64, P, P, P, P, P, P, P, I64, I64, INT, P, I64, I64, P, P, I64, P, P, P, SIZE, P)),
    "pysdm_b200_migration_merge": (
        INT, (P, I64, P, P, P, I64, P, P, P, I64, INT, P, I64, I64, P, P, P)),
    "pysdm_b200_collision_coalescence_breakup": (
        INT, (P, P, I64, P, I64, I64) + (P,) * 11 + (F64, P, INT, P, P)),
}



class FusedArgs(ctypes.Structure):
    """mirror of pysdm_b200_fused_args (include/pysdm_b200.h)"""
    _fields_ = ([(n, P) for n in (
        "idx_in", "idx_out", "cell_start", "cell_idx", "cell_id", "multiplicity", "attributes",
        "u01", "rand", "is_first_in_pair", "gamma", "norm_factor", "collision_rate",
        "collision_rate_deficit", "coalescence_rate", "healthy", "status")]
        + [(n, I64) for n in ("length", "n_sd", "n_cell", "n_attr", "mass_row",
                              "cells_per_batch")]
        + [(n, U64) for n in ("seed", "call_u01", "call_rand")]
        + [(n, F64) for n in ("timestep", "dv", "rho_w", "substeps", "kernel_a", "kernel_b")]
        + [("kernel", ctypes.c_int32), ("reserved", ctypes.c_int32), ("max_cell_size", I64),
           ("nf_slot", P), ("refresh_nf", I64), ("gate", P), ("epoch", I64),
           ("philox_round_keys", U32 * 20), ("shuffle_only", I64),
           ("radius", P), ("fall_velocity", P), ("berry_params", F64 * 13),
           ("berry_int_mask", I64), ("berry_unit", F64),
           ("adaptive", I64), ("dt_left", P), ("stats_n_substep", P), ("stats_dt_min", P),
           ("dt_min", F64), ("dt_max", F64), ("rank_to_cell", P), ("rank_info", P),
           ("adaptive_end", P), ("length_dev", P),
           ("dep_volume", P), ("dep_radius", P), ("dep_velocity", P), ("dep_table_b", P),
           ("dep_table_c", P), ("dep_radius_scale", F64), ("dep_radius_exponent", F64),
           ("dep_table_factor", F64), ("dep_radius_max", F64), ("dep_status", P),
           ("ws", P), ("ws_bytes", U64), ("sorted_length", I64), ("skip_coalescence", I64),
           ("prepass_done", I64)])


class BreakupArgs(ctypes.Structure):
    """mirror of pysdm_b200_breakup_args (include/pysdm_b200.h)"""
    _fields_ = ([(n, P) for n in (
        "multiplicity", "idx", "attributes", "gamma", "is_first_in_pair", "cell_id", "healthy",
        "coalescence_rate", "breakup_rate", "breakup_rate_deficit", "n_overflow")]
        + [(n, I64) for n in ("length", "n_sd", "n_attr", "mass_row")]
        + [("rand_proc", P), ("rand_frag", P)]
        + [(n, U64) for n in ("seed", "call_proc", "call_frag")]
        + [("volume", P), ("radius", P), ("fall_velocity", P),
           ("ec_kind", ctypes.c_int32), ("nf_kind", ctypes.c_int32),
           ("ec_const", F64), ("eb_const", F64), ("nf_n", F64), ("ec_params", F64 * 13),
           ("ec_int_mask", I64), ("ec_unit", F64), ("vmin", F64), ("nfmax", F64),
           ("has_nfmax", ctypes.c_int32), ("handle_all_breakups", ctypes.c_int32),
           ("max_multiplicity", F64)]
        + [(n, F64) for n in ("rho_w", "six_over_pi", "pi_sgm_w", "two_thirds", "nf_sc_factor",
                              "half_rho_w", "micro_joule")]
        + [("dep_volume", P), ("dep_radius", P), ("dep_velocity", P), ("dep_table_b", P),
           ("dep_table_c", P), ("dep_radius_scale", F64), ("dep_radius_exponent", F64),
           ("dep_table_factor", F64), ("dep_radius_max", F64), ("dep_status", P)])


KERNELS = {"Golovin": 0, "ConstantK": 1, "Linear": 2, "Geometric": 3, "Hydrodynamic": 4}
EW = {"add": 0, "sub": 1, "mul": 2, "div": 3, "div_if_not_zero": 4, "pow": 5, "axpy": 6,
      "floor": 7, "exp": 8, "abs": 9, "mod": 10}
PAIR_OPS = {"distance": 0, "max": 1, "min": 2, "sum": 3, "multiply": 4}


class CudaLibraryError(RuntimeError):
    pass


_lib = None


def load():
    """loads the shared library once; raises if it has not been built"""
    global _lib  # pylint: disable=global-statement
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise CudaLibraryError(
                f"{SO_PATH} is missing: build it with `python -m pysdm_b200.build` "
                "(pysdm_b200 has no CPU fallback)")
        lib = ctypes.CDLL(SO_PATH)
        for name, (restype, argtypes) in SIGNATURES.items():
            fun = getattr(lib, name)
            fun.restype = restype
            fun.argtypes = argtypes
        _lib = lib
    return _lib


def check(err, what=""):
    if err != 0:
        raise CudaLibraryError(f"CUDA error {err} in {what}")
