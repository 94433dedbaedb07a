"""`Random` of the B200 backend: counter-based Philox4x32-10 on the device.

Same call contract as the reference's generators (`Random(size, seed)`, callable on a
Storage; PySDM/backends/impl_common/random_common.py:6-10, impl_numba/random.py:9-19):
every call fills the storage with fresh uniforms in [0, 1).  Element e of call c under
seed s is a pure function of (s, c, e), independent of launch geometry.

`HostReplayRandom` reproduces the reference's Numba-backend stream exactly (numpy PCG64
drawn on the host, uploaded): with it a B200 run consumes the very same uniforms as a
`PySDM.backends.CPU` run with the same seed, which is what the parity tests use.
"""

import numpy as np
import torch

from . import _lib
from .recorder import RECORDER
from .storage import DEVICE, ptr, stream


class Random:  # pylint: disable=too-few-public-methods
    def __init__(self, size, seed):
        assert isinstance(size, int)
        assert isinstance(seed, (int, np.integer))
        self.size = size
        self.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        self.call = 0

    # the fill may be recorded instead of executed (pysdm_b200/recorder.py): when the step it
    # belongs to runs as a fused launch the kernel generates the same elements itself
    defer = False

    def __call__(self, storage):
        data, seed, call = storage.data, self.seed, self.call
        self.call += 1

        def run():
            _lib.check(_lib.load().pysdm_b200_philox_uniform(
                ptr(data), data.numel(), seed, call, stream()), "philox_uniform")

        if (self.defer and RECORDER.enabled and data.dim() == 1
                and RECORDER.defer("philox_fill", run, {"tensor": data, "seed": seed,
                                                        "call": call}, [data])):
            return
        if RECORDER.busy():
            RECORDER.sync([data])
        run()


class DeferredRandom(Random):  # pylint: disable=too-few-public-methods
    """`Random` of a B200 backend that records collision steps"""
    defer = True


class HostReplayRandom:  # pylint: disable=too-few-public-methods
    """numpy default_rng(seed).uniform on the host, copied to the device: bit-identical
    to the reference's CPU backend stream (impl_numba/random.py:17-18)"""

    def __init__(self, size, seed):
        self.size = size
        self.generator = np.random.default_rng(seed)

    def __call__(self, storage):
        host = self.generator.uniform(0, 1, storage.shape)
        if RECORDER.busy():
            RECORDER.sync([storage.data])
        storage.data.copy_(torch.from_numpy(host).to(DEVICE, non_blocking=False))
