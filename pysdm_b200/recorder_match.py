"""Recognises a recorded `Collision.step()` (pysdm_b200/recorder.py) and runs it as one fused
launch.  The sequence it looks for is the reference's, operation by operation
(PySDM/dynamics/collisions/collision.py:196-290, collision_kernels/*.py):

    [Random fill of pairs_rand, Random fill of rand]            random_generator_optimizer.py:37-48
    shuffle_local(idx, u01, cell_start)                         impl/particle_attributes.py:98-104
    find_pairs; sort_within_pair_by_attr(multiplicity)          collision.py:236-247
    <collision kernel writing kernel_temp>                      golovin.py / constantK.py /
                                                                geometric.py / impl/parameterized.py
    max_pair(prob, multiplicity); prob *= kernel_temp; normalize      collision.py:249-254
    prob /= substeps                                            collision.py:279
    compute_gamma(out = prob); collision_coalescence            collision.py:281-290, particulator.py:196-209

with the SAME storages threaded through (kernel_temp written by the kernel block is the one
multiplied into prob, prob is what compute_gamma reads and writes, ...).  Any deviation -> None,
and the recorder replays the record call by call.
"""

import torch

from . import methods as m

_PI = 3.141592653589793


def _same(a, b):
    return (isinstance(a, torch.Tensor) and isinstance(b, torch.Tensor)
            and a.data_ptr() == b.data_ptr() and a.numel() == b.numel())


class _Cursor:
    def __init__(self, ops):
        self.ops, self.i = ops, 0

    def peek(self, name):
        return self.i < len(self.ops) and self.ops[self.i].name == name

    def take(self, name, **expect):
        """next op if it has that name (and the given ew/pair operation), else None"""
        if not self.peek(name):
            return None
        args = self.ops[self.i].args
        for key, value in expect.items():
            if args.get(key) != value:
                return None
        self.i += 1
        return args


def _kernel_block(cur, backend):  # pylint: disable=too-many-return-statements,too-many-branches
    """-> dict(kernel=..., kernel_a, kernel_b, kt=<tensor written>, extra kwargs) or None"""
    first = cur.ops[cur.i] if cur.i < len(cur.ops) else None
    if first is None:
        return None
    if first.name == "fill":  # ConstantK: output.fill(a)
        a = cur.take("fill")
        return {"kernel": "ConstantK", "kernel_a": a["scalar"], "kernel_b": 0.0, "kt": a["out"]}
    if first.name == "lce":  # Parameterized: Hydrodynamic, Electric (impl/parameterized.py:13-30)
        lce = cur.take("lce")
        kt = lce["output"].data
        p2 = cur.take("ew", op="pow")
        mul_pi = cur.take("ew", op="mul")
        mx = cur.take("pair_op", op="max")
        p2b = cur.take("ew", op="pow")
        mul_tmp = cur.take("ew", op="mul")
        dist = cur.take("pair_op", op="distance")
        mul_tmp2 = cur.take("ew", op="mul")
        if None in (p2, mul_pi, mx, p2b, mul_tmp, dist, mul_tmp2):
            return None
        tmp = mx["out"].data
        ok = (_same(p2["out"], kt) and p2["scalar"] == 2 and p2["b"] is None
              and _same(mul_pi["out"], kt) and mul_pi["b"] is None and mul_pi["scalar"] == _PI
              and _same(mx["in"].data, lce["radii"].data)
              and _same(p2b["out"], tmp) and p2b["scalar"] == 2
              and _same(mul_tmp["out"], kt) and _same(mul_tmp["b"], tmp)
              and _same(dist["out"].data, tmp)
              and _same(mul_tmp2["out"], kt) and _same(mul_tmp2["b"], tmp))
        if not ok:
            return None
        return {"kernel": "Hydrodynamic", "kernel_a": _PI, "kernel_b": 0.0, "kt": kt,
                "radius": lce["radii"].data, "fall_velocity": dist["in"].data,
                "berry_params": lce["params"], "berry_unit": lce["unit"]}
    if first.name != "pair_op" or first.args["op"] != "sum":
        return None
    s = cur.take("pair_op", op="sum")
    kt, source = s["out"].data, s["in"].data
    if cur.peek("ew") and cur.ops[cur.i].args["op"] == "pow":  # Geometric (geometric.py:15-22)
        p2 = cur.take("ew", op="pow")
        mul_a = cur.take("ew", op="mul")
        dist = cur.take("pair_op", op="distance")
        mul_tmp = cur.take("ew", op="mul")
        if None in (p2, mul_a, dist, mul_tmp):
            return None
        tmp = dist["out"].data
        ok = (_same(p2["out"], kt) and p2["scalar"] == 2 and p2["b"] is None
              and _same(mul_a["out"], kt) and mul_a["b"] is None
              and _same(mul_tmp["out"], kt) and _same(mul_tmp["b"], tmp))
        if not ok:
            return None
        return {"kernel": "Geometric", "kernel_a": mul_a["scalar"], "kernel_b": 0.0, "kt": kt,
                "radius": source, "fall_velocity": dist["in"].data}
    # Golovin (golovin.py:14-16): sum_pair(volume); *= b -- the fused kernel derives the volume
    # from the water mass itself, so `source` must be a volume the backend computed from it
    mass = backend.volume_source(source)
    mul_b = cur.take("ew", op="mul")
    if mass is None or mul_b is None or not _same(mul_b["out"], kt) or mul_b["b"] is not None:
        return None
    return {"kernel": "Golovin", "kernel_a": 0.0, "kernel_b": mul_b["scalar"], "kt": kt,
            "mass": mass, "volume": source}


def fused_launch(rec, ops):  # pylint: disable=too-many-locals,too-many-return-statements,too-many-branches,too-many-statements
    """True when `ops` was one recognised step and has been executed as a fused launch"""
    cur = _Cursor(ops)
    fills = []
    while cur.peek("philox_fill"):
        fills.append(cur.take("philox_fill"))
    sh = cur.take("shuffle_local")
    fp = cur.take("find_pairs")
    sw = cur.take("sort_within_pair_by_attr")
    if None in (sh, fp, sw):
        return False
    backend = sh["backend"]
    kb = _kernel_block(cur, backend)
    if kb is None:
        return False
    mx = cur.take("pair_op", op="max")
    mul = cur.take("ew", op="mul")
    nz = cur.take("normalize")
    dv = cur.take("ew", op="div")
    cg = cur.take("compute_gamma")
    cc = cur.take(CHAIN_END_NAME)
    if None in (mx, mul, nz, dv, cg, cc) or cur.i != len(ops):
        return False
    idx, flags, mult = fp["idx"], fp["flags"], sw["attr"]
    prob = mx["out"].data
    cell_start = fp["cell_start"].data
    consistent = (
        _same(sh["idx"], idx.data) and _same(sh["cell_start"], cell_start)
        and sw["idx"] is idx and sw["flags"] is flags
        and mx["in"] is mult and mx["flags"] is flags and mx["idx"] is idx
        and _same(mul["out"], prob) and _same(mul["b"], kb["kt"])
        and _same(nz["prob"].data, prob) and _same(nz["cell_start"].data, cell_start)
        and nz["cell_idx"] is fp["cell_idx"] and _same(nz["cell_id"].data, fp["cell_id"].data)
        and _same(dv["out"], prob) and dv["b"] is None
        and _same(cg["prob"].data, prob) and _same(cg["out"].data, prob)
        and cg["multiplicity"] is mult and cg["flags"] is flags
        and _same(cc["gamma"].data, prob) and cc["multiplicity"] is mult and cc["idx"] is idx
        and cc["flags"] is flags and _same(cc["cell_id"].data, fp["cell_id"].data)
        and mult.idx is idx)
    facts = backend.sort_facts(cell_start)
    attributes = cc["attributes"].data
    n_attr, n_sd = attributes.shape
    u01, rand = sh["u01"], cg["rand"].data
    if (not consistent or facts is None or n_sd >= 2 ** 31 or facts.length != len(idx)
            or u01.numel() < n_sd or rand.numel() < n_sd // 2):
        return False
    kw = {k: kb[k] for k in ("radius", "fall_velocity", "berry_params", "berry_unit") if k in kb}
    mass_row = 0
    if "mass" in kb:  # which row of the extensive block the volume came from
        offset = kb["mass"].data_ptr() - attributes.data_ptr()
        if offset < 0 or offset % (8 * n_sd) or offset // (8 * n_sd) >= n_attr:
            return False
        mass_row = offset // (8 * n_sd)
        # the launch keeps the "volume" attribute current for the super-droplets it changes
        kw["dependents"] = {"volume": kb["volume"]}
    # (otherwise the row only matters for the derived attributes, which the stock path recomputes)
    # uniforms: generated in the kernel when both arrays come from recorded device fills
    philox = None
    if len(fills) == 2:
        f_u, f_r = fills
        if (_same(f_u["tensor"], u01) and _same(f_r["tensor"], rand)
                and f_u["seed"] == f_r["seed"]):
            philox = {"seed": f_u["seed"], "call_u01": f_u["call"], "call_rand": f_r["call"]}
    if philox is None:
        for f in fills:  # the arrays are read from memory: fill them now
            m.philox_uniform(f["tensor"], f["seed"], f["call"])
        philox = {"u01": u01, "rand": rand}
    # scratch of the fused step, per sorted population
    key = cell_start.data_ptr()
    scratch = rec.scratch.get(key)
    if scratch is None or scratch["nf_slot"].numel() != n_sd // 2:
        scratch = rec.scratch[key] = {
            "nf_slot": torch.empty(n_sd // 2, dtype=torch.float64, device=attributes.device),
            "status": torch.zeros(1, dtype=torch.int64, device=attributes.device), "built": None}
    built = (facts, facts.n_sorts, nz["timestep"], nz["dv"], fp["cell_idx"].data.data_ptr(),
             backend.version_of(fp["cell_idx"].data))
    err = m.fused_collision_step(
        idx_in=idx.data, idx_out=facts.tmp_idx.data, cell_start=cell_start,
        cell_idx=fp["cell_idx"].data, cell_id_=fp["cell_id"].data, multiplicity=mult.data,
        attributes=attributes, is_first_in_pair=flags.indicator.data, gamma=prob,
        norm_factor=nz["norm_factor"].data, collision_rate=cg["collision_rate"].data,
        collision_rate_deficit=cg["collision_rate_deficit"].data,
        coalescence_rate=cc["coalescence_rate"].data, healthy=cc["healthy"].data,
        status=scratch["status"], length=len(idx), sorted_length=facts.length,
        max_cell=facts.max_cell, kernel=kb["kernel"], kernel_a=kb["kernel_a"],
        kernel_b=kb["kernel_b"], timestep=nz["timestep"], dv=nz["dv"], rho_w=backend.rho_w,
        substeps=dv["scalar"], mass_row=mass_row, nf_slot=scratch["nf_slot"],
        refresh_nf=scratch["built"] != built, **philox, **kw)
    m.check(err, "fused_collision_step")
    scratch["built"] = built
    if facts.length < idx.data.numel():  # flagged positions beyond the sorted range stay
        facts.tmp_idx.data[facts.length:] = idx.data[facts.length:]
    idx.data, facts.tmp_idx.data = facts.tmp_idx.data, idx.data
    if "volume" in kb:
        rec.current[kb["volume"].data_ptr()] = (
            kb["mass"].data_ptr(), {kb["volume"].untyped_storage().data_ptr(),
                                    kb["mass"].untyped_storage().data_ptr()})
    else:
        rec.current = {}
    if "seed" in philox:  # the two arrays were never written: fill them if anybody asks
        for f in fills:
            rec.lazy[f["tensor"].untyped_storage().data_ptr()] = (f["tensor"], f["seed"], f["call"])
    return True


CHAIN_END_NAME = "collision_coalescence"
