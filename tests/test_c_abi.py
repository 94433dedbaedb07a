"""CPU: the drop-in boundary.  The C-ABI library loads without a GPU, exports every entry point
include/pysdm_b200.h declares, nothing in it carries a torch type, the ctypes table of
pysdm_b200/_lib.py mirrors the header one to one, and the product refuses to run without a
CUDA device (no CPU fallback)."""

import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pysdm_b200.h")


def declared_functions():
    text = open(HEADER, encoding="utf-8").read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)  # comments mention symbols too
    return sorted(set(re.findall(r"\b(pysdm_b200_\w+)\s*\(", text)))


@pytest.fixture(scope="module")
def library():
    from pysdm_b200 import build  # pylint: disable=import-outside-toplevel

    return ctypes.CDLL(build.build_extension())  # no-op when the in-tree .so is up to date


def test_header_is_plain_c():
    text = open(HEADER, encoding="utf-8").read()
    assert 'extern "C"' in text
    for banned in ("torch", "at::", "c10::", "std::", "template"):
        assert banned not in text, banned
    assert len(declared_functions()) >= 50


def test_every_declared_symbol_is_exported(library):
    missing = [name for name in declared_functions() if not hasattr(library, name)]
    assert not missing, missing


def test_ctypes_table_mirrors_header():
    from pysdm_b200 import _lib  # pylint: disable=import-outside-toplevel

    assert sorted(_lib.SIGNATURES) == declared_functions()
    text = re.sub(r"/\*.*?\*/", "", open(HEADER, encoding="utf-8").read(), flags=re.S)
    for name, (_, argtypes) in _lib.SIGNATURES.items():
        decl = re.search(r"\b" + name + r"\s*\((.*?)\)\s*;", text, flags=re.S).group(1)
        n_args = 0 if decl.strip() == "void" else decl.count(",") + 1
        assert n_args == len(argtypes), (name, n_args, len(argtypes))
    # the argument block of the fused step: same size on both sides of the boundary
    fields = re.search(r"typedef struct pysdm_b200_fused_args \{(.*?)\} pysdm_b200_fused_args;",
                       text, flags=re.S).group(1)
    assert ctypes.sizeof(_lib.FusedArgs) % 8 == 0 and "philox_round_keys[20]" in fields


def test_no_cpu_fallback():
    import torch  # pylint: disable=import-outside-toplevel

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from pysdm_b200 import methods  # pylint: disable=import-outside-toplevel
    from pysdm_b200._lib import CudaLibraryError  # pylint: disable=import-outside-toplevel
    from pysdm_b200.collision import DeviceCollision, Golovin  # pylint: disable=import-outside-toplevel

    with pytest.raises(CudaLibraryError):
        methods.identity_index(torch.zeros(4, dtype=torch.int64))
    with pytest.raises(RuntimeError):
        DeviceCollision(multiplicity=[1, 1], water_mass=[1.0, 1.0], dt=1.0, dv=1.0,
                        kernel=Golovin(1.0), seed=0)
