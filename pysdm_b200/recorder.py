"""Deferred execution of the collision step behind the STOCK dynamics.

`PySDM.dynamics.Coalescence.step()` (PySDM/dynamics/collisions/collision.py:196-234) issues its
work as ~15 separate backend calls: two `Random` fills, `shuffle_local`, `find_pairs`,
`sort_within_pair_by_attr`, the collision kernel's pair reductions and Storage operators,
`max_pair`, `*=`, `normalize`, `/=`, `compute_gamma`, `collision_coalescence`.  The B200 backend
does not execute those calls one by one: it RECORDS them (operation + the storages involved) and
looks at the record when `collision_coalescence` arrives.  If the record is exactly the
sequence of a step the fused kernel covers -- same storages threaded through it -- ONE launch
of `pysdm_b200_fused_collision_step` replaces it (csrc/fused*.cu: every storage the sequence
writes is written, bit for bit; only the dynamic's own scratch `kernel_temp` / `pair_tmp` keeps
its previous content).  Anything else is replayed call by call through the method-level
kernels, in order.

Consistency: every other backend method and every Storage operation first asks the recorder
whether it touches memory a recorded operation reads or writes; if so the record is replayed
first (`sync`).  Operations on unrelated memory -- e.g. the recalculation of the "volume"
attribute the orchestration triggers in the middle of the sequence -- run immediately.
Memory is identified by the base allocation of the tensors (views alias their base).

The two uniform arrays need not exist at all when the device generator feeds a fused launch:
the recorded `Random` fills then become "lazy" -- the kernel generates the same Philox
elements in place -- and are only materialised if something else reads those storages.
"""

import torch

from . import methods as m

CHAIN_END = "collision_coalescence"
MAX_QUEUE = 64


def base_ptr(tensor):
    return tensor.untyped_storage().data_ptr()


def tensors_of(obj, out, with_idx=True):
    """device tensors behind a backend-method argument (Storage / IndexedStorage / Index /
    PairIndicator / raw tensor).  `with_idx`: also the permutation an IndexedStorage carries
    (only the methods that read `x.idx` need it; the others get `idx` as an argument)"""
    if isinstance(obj, torch.Tensor):
        if obj.is_cuda:
            out.append(obj)
        return
    data = getattr(obj, "data", None)
    if isinstance(data, torch.Tensor) and data.is_cuda:
        out.append(data)
    for name in ("idx", "indicator") if with_idx else ("indicator",):
        sub = getattr(obj, name, None)
        if sub is not None and isinstance(getattr(sub, "data", None), torch.Tensor):
            out.append(sub.data)


class Op:  # pylint: disable=too-few-public-methods
    __slots__ = ("name", "run", "args", "ptrs", "written")

    def __init__(self, name, run, args, tensors, written):
        self.name, self.run, self.args = name, run, args
        self.ptrs = {base_ptr(t) for t in tensors}
        self.written = {base_ptr(t) for t in written}


class Recorder:  # pylint: disable=too-many-instance-attributes
    def __init__(self):
        self.enabled = True
        self.queue = []
        self.touched = set()   # base pointers the queued operations read or write
        self.written = set()   # ... write
        self.lazy = {}         # base pointer -> (tensor, seed, call): Random fills not yet done
        self.scratch = {}      # per cell_start tensor: nf_slot, status, what they were built for
        # data pointer of a "volume" tensor -> data pointer of the mass tensor it is current for
        # (a fused launch refreshed it for the super-droplets it changed, nothing wrote either
        # since): volume_of_water_mass has nothing to do then
        self.current = {}
        self.n_fused = self.n_replayed = 0

    # ---- recording ---------------------------------------------------------------------------
    def active(self):
        return self.enabled and bool(self.queue)

    def busy(self):
        """something is recorded, pending or being tracked: callers must `sync` what they touch"""
        return bool(self.queue or self.lazy or self.current)

    def defer(self, name, run, args, tensors, written=None):
        """records an operation; `run` executes it eagerly; `written`: the tensors among
        `tensors` it writes (default: all).  Returns True when recorded"""
        if not self.enabled:
            return False
        if self.lazy:
            if name == "philox_fill":  # a full overwrite: the pending fill is obsolete
                self.lazy.pop(base_ptr(tensors[0]), None)
            else:
                self._materialise(tensors)
        op = Op(name, run, args, tensors, tensors if written is None else written)
        self.queue.append(op)
        self.touched |= op.ptrs
        self.written |= op.written
        if name == CHAIN_END:
            self.resolve()
        elif len(self.queue) > MAX_QUEUE:
            self.flush()
        return True

    def sync(self, tensors, reads=()):
        """called before anything else writes (or may write) `tensors` and reads `reads`: a
        foreign write conflicts with whatever the record reads or writes, a foreign read only
        with what it writes"""
        if self.queue and (any(base_ptr(t) in self.touched for t in tensors)
                           or any(base_ptr(t) in self.written for t in reads)):
            self.flush()
        if self.current and tensors:
            written = {base_ptr(t) for t in tensors}
            self.current = {v: (mass, bases) for v, (mass, bases) in self.current.items()
                            if not bases & written}
        if self.lazy:
            self._materialise(tensors)
            self._materialise(reads)

    def _materialise(self, tensors):
        for t in tensors:
            entry = self.lazy.pop(base_ptr(t), None)
            if entry is not None:
                tensor, seed, call = entry
                m.philox_uniform(tensor, seed, call)

    def flush(self):
        """replays the record call by call"""
        ops, self.queue, self.touched, self.written = self.queue, [], set(), set()
        for op in ops:
            op.run()
        if ops:
            self.n_replayed += 1
            self.current = {}  # (the method-level kernels do not maintain derived attributes)

    def flush_all(self):
        """replays the record and performs every pending lazy fill (for code that is about to
        use device pointers directly)"""
        self.flush()
        self.current = {}
        for tensor, seed, call in self.lazy.values():
            m.philox_uniform(tensor, seed, call)
        self.lazy = {}

    def resolve(self):
        """the record ends with collision_coalescence: one fused launch, or the replay"""
        from .recorder_match import fused_launch  # pylint: disable=import-outside-toplevel

        ops, self.queue, self.touched, self.written = self.queue, [], set(), set()
        try:
            done = fused_launch(self, ops)
        except Exception:
            self.queue, self.touched, self.written = [], set(), set()
            raise
        if done:
            self.n_fused += 1
        else:
            for op in ops:
                op.run()
            self.n_replayed += 1
            self.current = {}


# one caller thread per process (SURVEY 8b "Threading"): one record
RECORDER = Recorder()
