"""Generates tests/golden/physics.npz by RUNNING THE REFERENCE (PySDM Numba backend,
Formulae(fastmath=False)): method-level input/output vectors for the rows of SURVEY 8(a)
beyond plain coalescence -- `**=` (signed pow), Gunn-Kinzer `interpolation`,
`terminal_velocity` (Rogers-Yau), `power_series`, `linear_collection_efficiency`, the
fragmentation functions with their limiters, `ll82_coalescence_check` and
`collision_coalescence_breakup` (both `break_up` and `break_up_while`).

TEST INFRASTRUCTURE (see gen_golden.py).  Usage:

    NUMBA_NUM_THREADS=1 python oracle/gen_golden_physics.py
"""

import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

import gen_golden as gg  # noqa: E402  (sets up the import path of the reference)
import numpy as np  # noqa: E402
from PySDM import Formulae  # noqa: E402
from PySDM.backends import CPU  # noqa: E402
from PySDM.dynamics.collisions.collision import DEFAULTS  # noqa: E402
from PySDM.dynamics.terminal_velocity import GunnKinzer1949, PowerSeries  # noqa: E402
from PySDM.physics import constants as const  # noqa: E402
from PySDM.physics import si  # noqa: E402

BERRY = (1, 1, -27, 1.65, -58, 1.9, 15, 1.13, 16.7, 1, 0.004, 4, 8)


def paired_population(rng, backend, types_, n_sd, n_cell):
    """a cell-sorted, shuffled population with pair flags, as Collision.step leaves it
    before the kernel is evaluated"""
    _, Index, IndexedStorage, PairIndicator, _ = types_
    S = backend.Storage
    cid, idx0, length = gg.random_cells(rng, n_sd, n_cell)
    idx = Index.from_ndarray(idx0.copy())
    idx.length = S.INT(length)
    cell_id = IndexedStorage.from_ndarray(idx, cid)
    cell_idx = Index.from_ndarray(np.arange(n_cell, dtype=np.int64))
    cell_start = S.from_ndarray(np.zeros(n_cell + 1, dtype=np.int64))
    backend.make_cell_caretaker(idx.shape, idx.dtype, n_cell + 1)(cell_id, cell_idx, cell_start,
                                                                  idx)
    idx.shuffle(S.from_ndarray(rng.uniform(0, 1, n_sd)), parts=cell_start)
    flag = PairIndicator(n_sd)
    flag.update(cell_start, cell_idx, cell_id)
    return cid, idx, cell_id, flag


def gen_physics():
    formulae = Formulae(fastmath=False, terminal_velocity="RogersYau")
    backend = CPU(formulae)
    types_ = (backend, gg.make_Index(backend), gg.make_IndexedStorage(backend),
              gg.make_PairIndicator(backend), gg.make_PairwiseStorage(backend))
    _, _, IndexedStorage, _, PairwiseStorage = types_
    S = backend.Storage
    rng = np.random.default_rng(20261105)
    g = {}

    def put(case, **kw):
        for k, v in kw.items():
            g[f"{case}/{k}"] = np.asarray(v)

    # ---- Storage.__ipow__ (storage_impl.py power): sign(x)|x|^e -----------------------
    x = np.concatenate([rng.exponential(1e-12, 400), -rng.exponential(3.0, 50),
                        rng.uniform(0, 5, 400), [0.0, 1.0, 8.0, 27.0, 1e-300, 1e300]])
    for tag, e in (("third", 1 / 3), ("two_thirds", 2 / 3), ("two", 2), ("minus_two", -2.0),
                   ("half", 0.5), ("p2_83", 2.83)):
        s = S.from_ndarray(x.copy())
        s **= e
        put(f"pow_{tag}", x=x, e=e, out=s.data.copy())

    # ---- radius from volume (attributes/physics/radius.py:15-17) -----------------------
    vol = rng.exponential(4 / 3 * np.pi * (30e-6) ** 3, 1000)
    r = S.empty(1000, dtype=float)
    r.product(S.from_ndarray(vol), 1 / const.PI_4_3)
    r **= 1 / 3
    put("radius", volume=vol, factor=1 / const.PI_4_3, out=r.data.copy())

    # ---- Gunn-Kinzer interpolation (dynamics/terminal_velocity/gunn_and_kinzer.py) -----
    gk = GunnKinzer1949(types.SimpleNamespace(backend=backend))
    radius = np.concatenate([rng.exponential(50e-6, 1500), rng.uniform(0, 5.9e-3, 500),
                             [0.0, -1e-6, 1e-5, 2e-5, 5.99e-3]])
    out = S.empty(len(radius), dtype=float)
    gk(out, S.from_ndarray(radius))
    put("interpolation", radius=radius, factor=gk.factor, a=gk.a.data, b=gk.b.data,
        out=out.data.copy())

    # ---- Rogers-Yau terminal velocity, power series ---------------------------------------
    radius = np.concatenate([rng.exponential(100e-6, 1000), [0.0, 35e-6, 600e-6, 34.9e-6, 1e-3]])
    out = S.empty(len(radius), dtype=float)
    backend.terminal_velocity(values=out.data, radius=radius)
    c = formulae.constants
    put("rogers_yau", radius=radius, out=out.data.copy(),
        params=(c.ROGERS_YAU_TERM_VEL_SMALL_K, c.ROGERS_YAU_TERM_VEL_MEDIUM_K,
                c.ROGERS_YAU_TERM_VEL_LARGE_K, c.ROGERS_YAU_TERM_VEL_SMALL_R_LIMIT,
                c.ROGERS_YAU_TERM_VEL_MEDIUM_R_LIMIT))
    ps = PowerSeries(types.SimpleNamespace(backend=backend),
                     prefactors=[2.0e-1, 3.0e-2], powers=[1 / 6, 1 / 3])
    out = S.empty(len(radius), dtype=float)
    ps(out, S.from_ndarray(radius))
    put("power_series", radius=radius, prefactors=ps.prefactors, powers=ps.powers,
        out=out.data.copy())

    # ---- linear_collection_efficiency (collisions_methods.py:743-782) ---------------------
    for tag, n_sd, n_cell in (("a", 61, 5), ("b", 2000, 1), ("c", 1500, 40)):
        cid, idx, _, flag = paired_population(rng, backend, types_, n_sd, n_cell)
        radii = rng.exponential(40e-6, n_sd) + 1e-6
        radii[rng.integers(0, n_sd, n_sd // 10)] = 20e-6  # equal radii: p == 1
        radii[rng.integers(0, n_sd, 3)] = 0.0  # p == 0
        radii[rng.integers(0, n_sd, n_sd // 10)] = 5e-4  # large drops
        out = PairwiseStorage.empty(n_sd // 2, dtype=float)
        backend.linear_collection_efficiency(
            params=BERRY, output=out, radii=IndexedStorage.from_ndarray(idx, radii),
            is_first_in_pair=flag, unit=const.si.um)
        print("lce", tag, flush=True)
        put(f"lce_{tag}", idx=idx.data.copy(), flag=flag.indicator.data.copy(), radii=radii,
            params=BERRY, unit=const.si.um, out=out.data.copy(), cell_id=cid)

    # ---- fragmentation functions (fragmentation_methods.py) ---------------------------------
    n = 3000
    x_plus_y = rng.exponential(4 / 3 * np.pi * (300e-6) ** 3, n)
    x_plus_y[:5] = 0.0
    # the fragmentation formulae are selected through Formulae(fragmentation_function=...)
    be_f = {name: CPU(Formulae(fastmath=False, fragmentation_function=name))
            for name in ("Feingold1988", "Straub2010Nf", "LowList1982Nf")}
    rand = rng.uniform(0, 1, n)
    rand[5:9] = (0.0, 1e-9, 0.999999, 0.5)
    vmin, nfmax = (0.01 * si.mm) ** 3 * np.pi / 6, 100

    def frag_out(**kw):
        out = {k: S.from_ndarray(np.full(n, np.nan)) for k in ("n_fragment", "frag_volume")}
        out["x_plus_y"] = S.from_ndarray(x_plus_y.copy())
        out["rand"] = S.from_ndarray(rand.copy())
        out.update(kw)
        return out

    for tag, lim in (("lim", dict(vmin=vmin, nfmax=nfmax)), ("nolim", dict(vmin=0.0, nfmax=None))):
        print("frag", tag, flush=True)
        a = frag_out()
        backend.exp_fragmentation(scale=1e-11, **a, **lim)
        put(f"frag_exp_{tag}", scale=1e-11, x_plus_y=x_plus_y, rand=rand,
            vmin=lim["vmin"], nfmax=np.nan if lim["nfmax"] is None else lim["nfmax"],
            n_fragment=a["n_fragment"].data, frag_volume=a["frag_volume"].data)
        a = frag_out()
        backend.gauss_fragmentation(mu=2e-11, sigma=1e-11, **a, **lim)
        put(f"frag_gauss_{tag}", mu=2e-11, sigma=1e-11,
            n_fragment=a["n_fragment"].data, frag_volume=a["frag_volume"].data)
        a = frag_out()
        be_f["Feingold1988"].feingold1988_fragmentation(scale=1e-11, fragtol=1e-3, **a, **lim)
        put(f"frag_feingold_{tag}", scale=1e-11, fragtol=1e-3,
            n_fragment=a["n_fragment"].data, frag_volume=a["frag_volume"].data)
        a = frag_out()
        probs = S.from_ndarray(np.full(n, np.nan))
        backend.slams_fragmentation(a["n_fragment"], a["frag_volume"], a["x_plus_y"], probs,
                                    a["rand"], lim["vmin"], lim["nfmax"])
        put(f"frag_slams_{tag}", probs=probs.data,
            n_fragment=a["n_fragment"].data, frag_volume=a["frag_volume"].data)

        # Straub 2010: physically plausible inputs, as breakup_fragmentations/straub2010.py
        # builds them
        rl = rng.uniform(0.2e-3, 3e-3, n)
        rs = rl * rng.uniform(0.05, 1.0, n)
        vl, vs = 4 / 3 * np.pi * rl ** 3, 4 / 3 * np.pi * rs ** 3
        du2 = rng.uniform(0.0, 8.0, n) ** 2
        sgm, rho_w = c.sgm_w, c.rho_w
        CKE = vl * vs / (vl + vs) * du2 * rho_w / 2
        Sc = (vl + vs) ** (2 / 3) * (np.pi * sgm * (6 / np.pi) ** (2 / 3))
        We = CKE / Sc
        CW = We * CKE / si.uJ
        CW[:20] = rng.uniform(0, 60, 20) * 1e6 / 1e6
        gam = rl / rs
        ds = 2 * rs
        a = frag_out(x_plus_y=S.from_ndarray(vl + vs))
        tmp = {k: S.from_ndarray(np.zeros(n)) for k in ("Nr1", "Nr2", "Nr3", "Nr4", "Nrt")}
        tmp["d34"] = S.from_ndarray(np.full(n, np.nan))
        be_f["Straub2010Nf"].straub_fragmentation(CW=S.from_ndarray(CW), gam=S.from_ndarray(gam),
                                     ds=S.from_ndarray(ds), v_max=S.from_ndarray(vl), **a, **tmp,
                                     **lim)
        put(f"frag_straub_{tag}", CW=CW, gam=gam, ds=ds, v_max=vl, x_plus_y=vl + vs,
            n_fragment=a["n_fragment"].data, frag_volume=a["frag_volume"].data,
            **{k: v.data for k, v in tmp.items()})

        # Low & List 1982, inputs as breakup_fragmentations/lowlist82.py builds them
        rl = rng.uniform(0.1e-3, 2.5e-3, n)
        rs = rl * rng.uniform(0.05, 1.0, n)
        rs[:4] = 0.0
        vl, vs = 4 / 3 * np.pi * rl ** 3, 4 / 3 * np.pi * rs ** 3
        dl, ds = 2 * rl, 2 * rs
        dcoal = ((vl + vs) / (np.pi / 6)) ** (1 / 3)
        Sc = (vl + vs) ** (2 / 3) * (np.pi * sgm * (6 / np.pi) ** (2 / 3))
        St = (ds ** 2 + dl ** 2) * np.pi * sgm
        CKE = vl * vs / (vl + vs) * du2 * rho_w / 2
        We, W2 = CKE / Sc, CKE / St
        a = frag_out(x_plus_y=S.from_ndarray(vl + vs))
        tmp = {k: S.from_ndarray(np.zeros(n)) for k in ("Rf", "Rs", "Rd")}
        be_f["LowList1982Nf"].ll82_fragmentation(CKE=S.from_ndarray(CKE), W=S.from_ndarray(We),
                                   W2=S.from_ndarray(W2), St=S.from_ndarray(St),
                                   ds=S.from_ndarray(ds), dl=S.from_ndarray(dl),
                                   dcoal=S.from_ndarray(dcoal), **a, **tmp, **lim)
        put(f"frag_ll82_{tag}", CKE=CKE, W=We, W2=W2, St=St, ds=ds, dl=dl, dcoal=dcoal,
            x_plus_y=vl + vs, rand_out=a["rand"].data,
            n_fragment=a["n_fragment"].data, frag_volume=a["frag_volume"].data,
            **{k: v.data for k, v in tmp.items()})
    Ec = S.from_ndarray(rng.uniform(0, 1, n))
    Ec_in = Ec.data.copy()
    backend.ll82_coalescence_check(Ec=Ec, dl=S.from_ndarray(dl))
    put("ll82_check", Ec_in=Ec_in, dl=dl, Ec_out=Ec.data)
    put("frag_common", x_plus_y=x_plus_y, rand=rand)

    # ---- collision_coalescence_breakup (collisions_methods.py:62-311, 479-520) ---------------
    for hab in (False, True):
        be2 = CPU(Formulae(fastmath=False, handle_all_breakups=hab))
        for tag, n_sd, n_cell, n_attr in (("a", 57, 6, 1), ("b", 1200, 1, 1), ("c", 900, 31, 3)):
            cid, idx, cell_id, flag = paired_population(rng, backend, types_, n_sd, n_cell)
            mult = rng.integers(1, 60, n_sd).astype(np.int64)
            mult[rng.integers(0, n_sd, n_sd // 8)] = 7  # equal multiplicities in some pairs
            if tag == "b" and not hab:
                # overflow of max_multiplicity (with handle_all_breakups the reference's
                # break_up_while never terminates on an overflowing pair: not generated)
                mult[:50] = 2 ** 40
            nn = IndexedStorage.from_ndarray(idx, mult.copy())
            backend.sort_within_pair_by_attr(idx, flag, nn)
            attrs = rng.exponential(1e-9, (n_attr, n_sd))
            a_st = IndexedStorage.from_ndarray(idx, attrs.copy())
            pmass = IndexedStorage.from_ndarray(idx, np.abs(attrs[0]).copy())
            p = n_sd // 2
            gamma = rng.integers(0, 4, p).astype(np.float64)
            if not hab:  # (large gamma overflows -> break_up_while would not terminate)
                gamma[rng.integers(0, p, p // 10)] = 25.0
            # slots without a pair carry gamma = 0 in a real step (prob = 0 there, quirk A6);
            # a nonzero gamma would make pair_indices hand out a droplet of the NEXT slot's pair
            fl = flag.indicator.data
            gamma[~(fl[0:2 * p:2] | fl[1:2 * p:2])] = 0.0
            rand_p = rng.uniform(0, 1, p)
            Ec_a = rng.uniform(0, 0.6, p)
            Eb_a = rng.uniform(0.5, 1.0, p)
            # fragment masses: some integer fractions of (m_j + m_k) (half-integer multiplicities
            # after the split), some arbitrary
            j = idx.data[0:2 * p:2]
            k = idx.data[1:2 * p:2]
            msum = np.abs(attrs[0, j]) + np.abs(attrs[0, k])
            frag = msum / rng.integers(1, 9, p)
            arb = rng.uniform(0, 1, p) < 0.5
            frag[arb] = (rng.exponential(3e-10, int(arb.sum())) + 1e-12 if not hab
                         else msum[arb] / rng.uniform(1.0, 8.0, int(arb.sum())))
            healthy = S.from_ndarray(np.full((1,), 1))
            rates = {k_: S.from_ndarray(np.zeros(n_cell, dtype=np.int64))
                     for k_ in ("coalescence_rate", "breakup_rate", "breakup_rate_deficit")}
            be2.collision_coalescence_breakup(
                multiplicity=nn, idx=idx, attributes=a_st, gamma=S.from_ndarray(gamma),
                rand=S.from_ndarray(rand_p), Ec=S.from_ndarray(Ec_a), Eb=S.from_ndarray(Eb_a),
                fragment_mass=S.from_ndarray(frag), healthy=healthy, cell_id=cell_id,
                is_first_in_pair=flag, warn_overflows=False, particle_mass=pmass,
                max_multiplicity=DEFAULTS.max_multiplicity, **rates)
            print("breakup", hab, tag, flush=True)
            put(f"breakup_{int(hab)}_{tag}", idx=idx.data.copy(), flag=flag.indicator.data.copy(),
                cell_id=cid, mult_in=mult, attrs_in=attrs, particle_mass=np.abs(attrs[0]),
                gamma=gamma, rand=rand_p, Ec=Ec_a, Eb=Eb_a, fragment_mass=frag,
                max_multiplicity=DEFAULTS.max_multiplicity, mult_out=nn.data, attrs_out=a_st.data,
                healthy=healthy.data, **{k_: v.data for k_, v in rates.items()})

    os.makedirs(gg.OUT, exist_ok=True)
    np.savez_compressed(os.path.join(gg.OUT, "physics.npz"), **g)
    print("physics.npz:", len(g), "arrays")


if __name__ == "__main__":
    gen_physics()
