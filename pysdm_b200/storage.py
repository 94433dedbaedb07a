"""`Storage` of the B200 backend: a typed view of device memory (a torch CUDA tensor is
used purely as the allocation/slicing handle) with the in-place operator set the
reference's storage defines (PySDM/backends/impl_numba/storage.py:16-213 and
impl_common/storage_utils.py:16-79).  Arithmetic goes through our own kernels
(csrc/storage_ops.cu); torch is only used to allocate, slice, fill and copy.
"""

import ctypes
from collections import namedtuple

import numpy as np
import torch

from . import _lib
from .recorder import RECORDER

try:  # when PySDM is importable the reference's own base classes are used (drop-in)
    from PySDM.backends.impl_common.storage_utils import StorageBase, StorageSignature
except ImportError:  # stand-alone: same fields, same dunder behaviour
    class StorageSignature(namedtuple("StorageSignature", ("data", "shape", "dtype"))):
        __slots__ = ()

    class StorageBase:
        def __init__(self, signature):
            self.data = signature.data
            self.shape = ((signature.shape,) if isinstance(signature.shape, (int, np.integer))
                          else signature.shape)
            self.dtype = signature.dtype
            self.backend = None

        def __len__(self):
            return self.shape[0]

        @staticmethod
        def _use_inplace(symbol):
            raise TypeError(f"Use {symbol}=")

        def __pow__(self, other):
            self._use_inplace("**")

        def __mod__(self, other):
            self._use_inplace("%")

        def __truediv__(self, other):
            self._use_inplace("/")

        def __mul__(self, other):
            self._use_inplace("*")

        def __sub__(self, other):
            self._use_inplace("-")

        def __add__(self, other):
            self._use_inplace("+")


DEVICE = "cuda"
_TORCH = {np.float64: torch.float64, np.int64: torch.int64, np.bool_: torch.bool}


def stream():
    from .methods import _stream  # pylint: disable=import-outside-toplevel

    return _stream()


def ptr(tensor):
    """raw device pointer of a contiguous tensor (or None)"""
    if tensor is None:
        return None
    assert tensor.is_contiguous(), "non-contiguous device view"
    return tensor.data_ptr()


class _Scratch:
    """per-process device scratch + a pinned two-word mailbox for host results"""

    def __init__(self):
        self.buffers = {}
        self.mailbox = None

    def device(self, key, nbytes):
        buf = self.buffers.get(key)
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty(int(nbytes) + 256, dtype=torch.uint8, device=DEVICE)
            self.buffers[key] = buf
        return buf

    def host(self):
        if self.mailbox is None:
            self.mailbox = torch.zeros(2, dtype=torch.int64).pin_memory()
        return self.mailbox


SCRATCH = _Scratch()


class Storage(StorageBase):
    FLOAT = np.float64
    INT = np.int64
    BOOL = np.bool_

    # ---- construction -------------------------------------------------------------
    @staticmethod
    def _normalize_dtype(dtype):
        if dtype in (float, Storage.FLOAT):
            return Storage.FLOAT
        if dtype in (int, Storage.INT):
            return Storage.INT
        if dtype in (bool, Storage.BOOL):
            return Storage.BOOL
        raise NotImplementedError(dtype)

    @staticmethod
    def _get_empty_data(shape, dtype):
        dtype = Storage._normalize_dtype(dtype)
        size = (shape,) if isinstance(shape, (int, np.integer)) else tuple(shape)
        # same initial content as the reference: NaN / -1 / True
        fill = {Storage.FLOAT: float("nan"), Storage.INT: -1, Storage.BOOL: True}[dtype]
        data = torch.full(size, fill, dtype=_TORCH[dtype], device=DEVICE)
        return StorageSignature(data, shape, dtype)

    @staticmethod
    def empty(shape, dtype):
        return Storage(Storage._get_empty_data(shape, dtype))

    @staticmethod
    def _get_data_from_ndarray(array):
        array = np.asarray(array)
        kind = str(array.dtype)
        if kind.startswith("int"):
            dtype = Storage.INT
        elif kind.startswith("float"):
            dtype = Storage.FLOAT
        elif kind.startswith("bool"):
            dtype = Storage.BOOL
        else:
            raise NotImplementedError(kind)
        host = np.ascontiguousarray(array.astype(dtype))
        data = torch.from_numpy(host).to(DEVICE)
        return StorageSignature(data, array.shape, dtype)

    @staticmethod
    def from_ndarray(array):
        return Storage(Storage._get_data_from_ndarray(array))

    # ---- views & element access -----------------------------------------------------
    def __getitem__(self, item):
        dim = len(self.shape)
        if isinstance(item, slice):
            if (item.step or 1) != 1:
                raise NotImplementedError("step != 1")
            start = item.start or 0
            if dim == 1:
                stop = item.stop or len(self)
                shape = (stop - start,)
            elif dim == 2:
                stop = item.stop or self.shape[0]
                shape = (stop - start, self.shape[1])
            else:
                raise NotImplementedError("Only 2 or less dimensions array is supported.")
            if stop > self.data.shape[0]:
                raise IndexError(f"requested a slice ({start}:{stop}) of Storage"
                                 f" with first dim of length {self.data.shape[0]}")
            return Storage(StorageSignature(self.data[item], shape, self.dtype))
        if isinstance(item, tuple) and dim == 2 and isinstance(item[1], slice):
            return Storage(StorageSignature(self.data[item[0]], (*self.shape[1:],), self.dtype))
        self._peek()
        return self.data[item].item()  # host scalar (device sync)

    def __setitem__(self, key, value):
        if hasattr(value, "data") and isinstance(value.data, torch.Tensor):
            value = value.data
        elif isinstance(value, np.ndarray):
            value = torch.from_numpy(np.ascontiguousarray(value)).to(DEVICE)
        self._touch(value)
        self.data[key] = value
        return self

    # ---- arithmetic (own kernels) -----------------------------------------------------
    def _touch(self, *others):
        """before this storage's memory is written (and `others` read) outside the record of a
        collision step (pysdm_b200/recorder.py)"""
        if RECORDER.busy():
            RECORDER.sync([self.data], [o for o in others if isinstance(o, torch.Tensor)])

    def _peek(self):
        """before this storage's memory is read"""
        if RECORDER.busy():
            RECORDER.sync((), [self.data])

    def _ew(self, op, a=None, b=None, scalar=0.0, out=None):
        assert self.dtype is Storage.FLOAT, "elementwise arithmetic is defined for fp64 storages"
        out = self.data if out is None else out

        def run():
            _lib.check(_lib.load().pysdm_b200_elementwise(
                _lib.EW[op], ptr(out), ptr(a), ptr(b), float(scalar), out.numel(), stream()), op)

        if RECORDER.busy():
            tensors = [t for t in (out, a, b) if t is not None]
            # in-place `*= += **= /=` on storages a recorded collision step has written are part
            # of the record (e.g. `output *= b` after sum_pair); unrelated ones (the `**= 1/3` of
            # the radius attribute, recalculated in the middle of a step) run at once
            if (RECORDER.queue and RECORDER.enabled and a is out
                    and op in ("mul", "add", "pow", "div")
                    and any(t.untyped_storage().data_ptr() in RECORDER.touched for t in tensors)
                    and RECORDER.defer("ew", run, {"op": op, "out": out, "a": a, "b": b,
                                                   "scalar": float(scalar)}, tensors, [out])):
                return self
            RECORDER.sync([out], [t for t in (a, b) if t is not None and t is not out])
        run()
        return self

    @staticmethod
    def _operand(other):
        """-> (tensor or None, scalar)"""
        if hasattr(other, "data") and isinstance(other.data, torch.Tensor):
            return other.data, 0.0
        if isinstance(other, torch.Tensor):
            return other, 0.0
        return None, float(other)

    def __iadd__(self, other):
        if (isinstance(other, tuple) and len(other) == 3 and isinstance(other[0], float)
                and other[1] == "*" and isinstance(other[2], Storage)):
            return self._ew("axpy", self.data, other[2].data, other[0])
        if self.dtype is not Storage.FLOAT:
            tensor, scalar = self._operand(other)
            self._touch(tensor)
            self.data += tensor if tensor is not None else int(scalar)
            return self
        tensor, scalar = self._operand(other)
        return self._ew("add", self.data, tensor, scalar)

    def __isub__(self, other):
        if self.dtype is Storage.FLOAT and getattr(other, "dtype", None) is Storage.INT:
            # position_in_cell -= floor_of_position (dynamics/displacement.py:150)
            self._touch(other.data)
            _lib.check(_lib.load().pysdm_b200_sub_i64(ptr(self.data), ptr(other.data),
                                                     self.data.numel(), stream()), "sub_int")
            return self
        return self._ew("sub", self.data, other.data)

    def __imul__(self, other):
        tensor, scalar = self._operand(other)
        return self._ew("mul", self.data, tensor, scalar)

    def __itruediv__(self, other):
        tensor, scalar = self._operand(other)
        return self._ew("div", self.data, tensor, scalar)

    def __imod__(self, other):
        n_rows, n_cols = (self.shape if len(self.shape) == 2 else (1, self.shape[0]))
        self._touch(other.data)
        if self.dtype is Storage.INT:  # cell_origin %= grid (dynamics/displacement.py:152-153)
            _lib.check(_lib.load().pysdm_b200_row_modulo_i64(
                ptr(self.data), ptr(other.data), n_rows, n_cols, stream()), "row_modulo")
            return self
        _lib.check(_lib.load().pysdm_b200_row_modulo(ptr(self.data), ptr(other.data), n_rows,
                                                    n_cols, stream()), "row_modulo")
        return self

    def __ipow__(self, other):
        return self._ew("pow", self.data, None, float(other))

    def floor(self, other=None):
        if self.dtype is Storage.INT:  # floor_of_position.floor(position_in_cell), :147-148
            self._touch(other.data)
            _lib.check(_lib.load().pysdm_b200_floor_to_i64(
                ptr(self.data), ptr(other.data), self.data.numel(), stream()), "floor")
            return self
        return self._ew("floor", None, self.data if other is None else other.data)

    def product(self, multiplicand, multiplier):
        tensor, scalar = self._operand(multiplier)
        return self._ew("mul", multiplicand.data, tensor, scalar)

    def ratio(self, dividend, divisor):
        return self._ew("div", dividend.data, divisor.data)

    def divide_if_not_zero(self, divisor):
        return self._ew("div_if_not_zero", self.data, divisor.data)

    def sum(self, arg_a, arg_b):
        return self._ew("add", arg_a.data, arg_b.data)

    def exp(self):
        self._ew("exp", None, self.data)

    def abs(self):
        self._ew("abs", None, self.data)

    # ---- reductions (host result => device sync) ---------------------------------------
    def _minmax(self, is_max):
        self._peek()
        lib = _lib.load()
        scratch = SCRATCH.device("reduce", lib.pysdm_b200_reduce_scratch_bytes())
        box = SCRATCH.host()
        data = self.data.reshape(-1)
        if self.dtype is Storage.FLOAT:
            fun = lib.pysdm_b200_minmax_f64
        elif self.dtype is Storage.INT:
            fun = lib.pysdm_b200_minmax_i64
        else:
            raise NotImplementedError()
        _lib.check(fun(ptr(data), data.numel(), int(is_max), box.data_ptr(), ptr(scratch), stream()),
                   "minmax")
        if self.dtype is Storage.FLOAT:
            if int(box[1]) != 0:
                return float("nan")  # numpy's amin/amax propagate NaN
            return float(box[:1].view(torch.float64)[0])
        return int(box[0])

    def amin(self):
        return self._minmax(False)

    def amax(self):
        return self._minmax(True)

    def all(self):
        self._peek()
        return bool(torch.all(self.data != 0).item())

    def __bool__(self):
        if len(self) == 1:
            self._peek()
            return bool(self.data[0].item() != 0)
        raise NotImplementedError("Logic value of array is ambiguous.")

    # ---- memory plumbing ---------------------------------------------------------------
    def detach(self):
        self._touch()
        if self.data._base is not None:  # pylint: disable=protected-access
            self.data = self.data.clone()

    def download(self, target, reshape=False, non_blocking=False):
        self._peek()
        if isinstance(target, torch.Tensor):  # straight into (pinned) host memory
            # non_blocking: the caller synchronises the stream before it reads `target`
            target.copy_(self.data.reshape(target.shape), non_blocking=non_blocking)
            return
        host = self.data.cpu().numpy()
        if reshape:
            host = host.reshape(target.shape)
        np.copyto(target, host, casting="safe")

    def to_ndarray(self):
        self._peek()
        return self.data.cpu().numpy().copy()

    def upload(self, data):
        self._touch()
        if isinstance(data, torch.Tensor):  # (pinned) host tensor of the storage's own dtype
            self.data.copy_(data.reshape(self.data.shape), non_blocking=True)
            return
        host = np.ascontiguousarray(np.asarray(data).astype(self.dtype, casting="safe"))
        self.data.copy_(torch.from_numpy(host).reshape(self.data.shape))

    def ravel(self, other):
        src = other.data if isinstance(other, Storage) else torch.from_numpy(
            np.ascontiguousarray(other)).to(DEVICE)
        self._touch(src)
        self.data[:] = src.reshape(-1)

    def urand(self, generator):
        generator(self)

    def fill(self, other):
        if (RECORDER.queue and RECORDER.enabled and self.dtype is Storage.FLOAT
                and isinstance(other, (int, float)) and not isinstance(other, bool)):
            out, value = self.data, float(other)  # `output.fill(a)` of a recorded collision step

            def run():
                out[:] = value

            if RECORDER.defer("fill", run, {"out": out, "scalar": value}, [out]):
                return
        self._touch(other.data if isinstance(other, Storage) else None)
        if isinstance(other, Storage):
            self.data[:] = other.data
        elif isinstance(other, np.ndarray):
            self.upload(other)
        elif other is None:  # numpy semantics of `float_array[:] = None`
            self.data[:] = float("nan")
        else:
            self.data[:] = other

    # ---- numpy interoperability -----------------------------------------------------------
    # The reference's orchestration applies numpy ufuncs to Storage objects in one place
    # (`output.fill(np.exp(We))`, dynamics/collisions/coalescence_efficiencies/straub2010.py);
    # with the Numba backend numpy walks the Storage through the sequence protocol.  Here
    # np.exp stays on the device (own kernel) and returns a Storage; any other ufunc gets a
    # host copy through __array__.
    def __array_ufunc__(self, ufunc, method, *inputs, **kwargs):
        if ufunc is np.exp and method == "__call__" and len(inputs) == 1 and not kwargs:
            out = Storage.empty(self.shape, self.dtype)
            out._ew("exp", None, self.data)  # pylint: disable=protected-access
            return out
        arrays = [i.to_ndarray() if isinstance(i, Storage) else i for i in inputs]
        return getattr(ufunc, method)(*arrays, **kwargs)

    def __array__(self, dtype=None, copy=None):  # pylint: disable=unused-argument
        host = self.to_ndarray()
        return host if dtype is None else host.astype(dtype)


def host_int_result():
    """pinned int64 mailbox pointer for C-ABI calls that return a host integer"""
    box = SCRATCH.host()
    return box, ctypes.c_void_p(box.data_ptr())
