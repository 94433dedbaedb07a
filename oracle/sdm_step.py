"""CPU restatement of the reference's collision time step on top of the C oracle.

TEST INFRASTRUCTURE, NOT PRODUCT CODE (see oracle/sdm_oracle.c header).  This
is the checker the CUDA path is compared with and the thing bench.py times as
`cpu_baseline` (kind "port").

Follows, call by call:
  Collision.__call__ / step / compute_gamma   PySDM/dynamics/collisions/collision.py:174-290
  ParticleAttributes (sort, permutation, sanitize, working length)
                                              PySDM/impl/particle_attributes.py:42-111
  RandomGeneratorOptimizer                    PySDM/dynamics/impl/random_generator_optimizer.py:19-48
  Particulator.collision_coalescence_breakup  PySDM/particulator.py:157-213
with numpy arrays as state and oracle/sdm_oracle.py as the "backend".
"""

import math

import numpy as np

from . import sdm_oracle as orc


class NumpyRandom:
    """the reference's Random: numpy PCG64, stream advances per call
    (PySDM/backends/impl_numba/random.py:9-19)"""

    def __init__(self, size, seed):
        self.size = size
        self.generator = np.random.default_rng(seed)

    def __call__(self, array):
        array[:] = self.generator.uniform(0, 1, array.shape)


class Golovin:
    """collision_kernels/golovin.py:14-16"""

    def __init__(self, b):
        self.b = b

    def __call__(self, box, output, flag):
        box.be.pair_op("sum", output, box.volume(), flag, box.idx, box.length)
        output *= self.b


class ConstantK:
    """collision_kernels/constantK.py:11-12"""

    def __init__(self, a):
        self.a = a

    def __call__(self, box, output, flag):
        output.fill(self.a)


class Linear:
    """collision_kernels/linear.py:12-15 as intended (the reference spells the pair sum
    `output.sum_pair`, which PairwiseStorage does not have -- its own TODO #744; Golovin,
    golovin.py:14-16, is the same sequence without the `+= a`)"""

    def __init__(self, a, b):
        self.a, self.b = a, b

    def __call__(self, box, output, flag):
        box.be.pair_op("sum", output, box.volume(), flag, box.idx, box.length)
        output *= self.b
        output += self.a


class Geometric:
    """collision_kernels/geometric.py:15-22 on the attributes the orchestration derives from the
    water mass before the kernel reads them: volume (attributes/physics/volume.py:17), radius
    (radius.py:15-17: product with 1 / PI_4_3, then `**= 1/3`), fall velocity (Gunn-Kinzer
    `interpolation`, dynamics/terminal_velocity/gunn_and_kinzer.py:128-133); `table` = (a, b,
    factor, maximum_radius) of the reference's GunnKinzer1949 object"""
    PI = 3.141592653589793
    PI_4_3 = PI * 4 / 3

    def __init__(self, collection_efficiency, table):
        self.collection_efficiency = collection_efficiency
        self.table = table
        self.radius = self.velocity = self.pair_tmp = None

    def __call__(self, box, output, flag):
        be = box.be
        if self.radius is None:
            self.radius, self.velocity = np.empty(box.n_sd), np.empty(box.n_sd)
            self.pair_tmp = np.empty(box.n_sd // 2)
        be.product_scalar(self.radius, box.volume(), 1 / self.PI_4_3)
        be.power(self.radius, 1 / 3)
        a, b, factor, r_max = self.table
        assert not self.radius.max() > r_max
        be.interpolation(self.velocity, self.radius, factor, a, b)
        be.pair_op("sum", output, self.radius, flag, box.idx, box.length)
        be.power(output, 2)
        output *= self.PI * self.collection_efficiency
        be.pair_op("distance", self.pair_tmp, self.velocity, flag, box.idx, box.length)
        output *= self.pair_tmp


class OracleCollision:  # pylint: disable=too-many-instance-attributes
    """state + time step of `Coalescence(collision_kernel=...)` in a Box or on a mesh"""

    def __init__(self, *, multiplicity, water_mass, cell_id=None, n_cell=1, dt, dv, kernel,
                 seed, adaptive=False, substeps=1, dt_coal_range=(0.1, 100.0), croupier="local",
                 optimized_random=False, rho_w=1000.0, extra_attributes=None, rnd=None, be=None):
        assert substeps == 1 or adaptive is False
        # `be`: module with the signatures of oracle/sdm_oracle.py (tests plug the CUDA
        # adapter in here to drive the device kernels through the very same step logic)
        self.be = be or orc
        self.kernel_be = self.be
        self.n_sd = len(multiplicity)
        self.n_cell = n_cell
        self.dt, self.dv, self.kernel, self.rho_w = dt, dv, kernel, rho_w
        self.adaptive, self.substeps, self.croupier = adaptive, substeps, croupier
        self.dt_coal_range = (dt_coal_range[0], min(dt_coal_range[1], dt))
        self.multiplicity = np.ascontiguousarray(multiplicity, dtype=np.int64).copy()
        rows = [np.asarray(water_mass, dtype=np.float64)]
        rows += [np.asarray(a, dtype=np.float64) for a in (extra_attributes or [])]
        self.attributes = np.ascontiguousarray(np.stack(rows))
        self.cell_id = (np.zeros(self.n_sd, np.int64) if cell_id is None
                        else np.ascontiguousarray(cell_id, dtype=np.int64).copy())
        # ParticleAttributes.__init__
        self.idx = np.arange(self.n_sd, dtype=np.int64)
        self.tmp_idx = np.full(self.n_sd, -1, dtype=np.int64)
        self.length = self.n_sd
        self.valid_n_sd = self.n_sd
        self.cell_idx = np.arange(n_cell, dtype=np.int64)
        self.cell_start = np.zeros(n_cell + 1, dtype=np.int64)
        self.sorted = False
        self.healthy = np.ones(1, dtype=np.int64)
        # Collision.register
        n_pairs = self.n_sd // 2
        self.kernel_temp = np.full(n_pairs, np.nan)
        self.gamma = np.full(n_pairs, np.nan)
        self.flag = np.ones(self.n_sd, dtype=np.bool_)
        self.norm_factor = np.full(n_cell, np.nan)
        self.dt_left = np.full(n_cell, np.nan)
        self.stats_n_substep = np.full(n_cell, 0 if adaptive else substeps, dtype=np.int64)
        self.stats_dt_min = np.full(n_cell, np.nan)
        self.collision_rate = np.zeros(n_cell, dtype=np.int64)
        self.collision_rate_deficit = np.zeros(n_cell, dtype=np.int64)
        self.coalescence_rate = np.zeros(n_cell, dtype=np.int64)
        # RandomGeneratorOptimizer.register
        self.optimized_random = optimized_random
        shift = math.ceil(dt / dt_coal_range[0]) if optimized_random else 0
        self.pairs_rand = np.full(self.n_sd + shift, np.nan)
        self.rand = np.full(n_pairs, np.nan)
        self.rnd = rnd or NumpyRandom(self.n_sd + shift, seed)
        self.substep = 0
        self._volume = np.full(self.n_sd, np.nan)
        self.dt_min_warning = False
        # Builder.build: zero multiplicities are removed up front
        if (self.multiplicity == 0).any():
            self.healthy[0] = 0
            self.sanitize()

    # ---- attributes -------------------------------------------------------------
    def volume(self):
        self.be.volume_of_water_mass(self._volume, self.attributes[0], self.rho_w)
        return self._volume

    def sort_by_cell_id(self):
        self.be.counting_sort_by_cell_id(self.tmp_idx, self.idx, self.cell_id, self.cell_idx,
                                     self.length, self.cell_start)
        self.idx, self.tmp_idx = self.tmp_idx, self.idx
        self.sorted = True

    def get_cell_start(self):
        if not self.sorted:
            self.sort_by_cell_id()
        return self.cell_start

    def sanitize(self):
        if not self.healthy[0]:
            self.length = self.valid_n_sd
            self.length = self.be.remove_zero_n_or_flagged(self.multiplicity, self.idx, self.length)
            self.valid_n_sd = self.length
            self.healthy[0] = 1
            self.sorted = False

    def permutation(self, u01, local):
        if local:
            cell_start = self.get_cell_start()  # may swap self.idx <-> self.tmp_idx
            self.be.shuffle_local(self.idx, u01, cell_start)
        else:
            self.be.shuffle_global(self.idx, self.length, u01)
            self.sorted = False

    # ---- random numbers -----------------------------------------------------------
    def get_random_arrays(self):
        if self.optimized_random:
            shift = self.substep
            if self.substep == 0:
                self.rnd(self.pairs_rand)
                self.rnd(self.rand)
        else:
            shift = 0
            self.rnd(self.pairs_rand)
            self.rnd(self.rand)
        self.substep += 1
        return self.pairs_rand[shift:self.n_sd + shift], self.rand

    # ---- the dynamic ---------------------------------------------------------------
    def __call__(self):
        if not self.adaptive:
            for _ in range(self.substeps):
                self.step()
        else:
            self.dt_left[:] = self.dt
            while self.length != 0:
                self.be.sort_by_key(self.cell_idx, self.dt_left)
                self.step()
                end = self.be.adaptive_sdm_end(self.dt_left, self.get_cell_start())
                assert end <= self.length
                self.length = end
            self.length = self.valid_n_sd
            self.be.identity_index(self.cell_idx)
            self.sort_by_cell_id()
        self.substep = 0

    def step(self):
        pairs_rand, rand = self.get_random_arrays()
        # toss_candidate_pairs_and_sort_within_pair_by_multiplicity
        self.permutation(pairs_rand, self.croupier == "local")
        cell_start = self.get_cell_start()
        self.be.find_pairs(cell_start, self.flag, self.cell_id, self.cell_idx, self.idx, self.length)
        self.be.sort_within_pair_by_attr(self.idx, self.length, self.flag, self.multiplicity)
        # compute_probabilities_of_collision
        prob = self.gamma
        self.kernel(self, self.kernel_temp, self.flag)
        self.be.pair_op("max", prob, self.multiplicity, self.flag, self.idx, self.length)
        prob *= self.kernel_temp
        self.be.normalize(prob, self.cell_id, self.cell_idx, self.get_cell_start(), self.norm_factor,
                      self.dt, self.dv)
        # compute_gamma
        if self.adaptive:
            self.be.scale_prob_for_adaptive_sdm_gamma(prob, self.idx, self.length, self.multiplicity,
                                                  self.cell_id, self.dt_left, self.dt,
                                                  self.dt_coal_range, self.flag,
                                                  self.stats_n_substep, self.stats_dt_min)
            if np.amin(self.stats_dt_min) == self.dt_coal_range[0]:
                self.dt_min_warning = True
        else:
            prob /= self.substeps
        self.be.compute_gamma(prob, rand, self.idx, self.length, self.multiplicity, self.cell_id,
                          self.collision_rate_deficit, self.collision_rate, self.flag, self.gamma)
        # collision_coalescence_breakup
        self.be.collision_coalescence(self.multiplicity, self.idx, self.length, self.attributes,
                                  self.gamma, self.healthy, self.cell_id, self.coalescence_rate,
                                  self.flag)
        self.sanitize()
