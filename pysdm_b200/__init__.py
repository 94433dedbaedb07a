"""pysdm_b200 -- the Super-Droplet-Method collision / coalescence / breakup step of PySDM
as hand-written CUDA for NVIDIA B200 (sm_100a), behind PySDM's backend interface.

    from pysdm_b200 import B200        # backend class: Builder(backend=B200(), ...)

The CUDA library (libpysdm_b200.so, built in-tree by `python -m pysdm_b200.build`) is
mandatory: there is no CPU fallback.
"""

__all__ = ("B200",)


def __getattr__(name):  # lazy: importing the package must not need torch/CUDA
    if name == "B200":
        from .backend import B200  # pylint: disable=import-outside-toplevel

        return B200
    raise AttributeError(name)
