"""pysdm_b200 -- the Super-Droplet-Method collision / coalescence / breakup step of PySDM
as hand-written CUDA for NVIDIA B200 (sm_100a), behind PySDM's backend interface.

    from pysdm_b200 import B200        # backend class: Builder(backend=B200(), ...)

The CUDA library (libpysdm_b200.so, built in-tree by `python -m pysdm_b200.build`) is
mandatory: there is no CPU fallback.
"""

__all__ = ("B200", "FusedCoalescence", "FusedCollision")


def __getattr__(name):  # lazy: importing the package must not need torch/CUDA (or PySDM)
    if name == "B200":
        from .backend import B200  # pylint: disable=import-outside-toplevel

        return B200
    if name in ("FusedCoalescence", "FusedCollision"):  # PySDM's dynamics on the fused kernels
        from .dynamics import make_fused_coalescence  # pylint: disable=import-outside-toplevel

        cls = make_fused_coalescence(name)
        globals()[name] = cls
        return cls
    raise AttributeError(name)
