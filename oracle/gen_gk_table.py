"""Writes pysdm_b200/data/gunn_kinzer_1949.npz: the terminal-velocity interpolation table the
reference builds at start-up (PySDM/dynamics/terminal_velocity/gunn_and_kinzer.py:15-125: SciPy
`Rbf` through Gunn & Kinzer's Table 2 + the small-radius formula), taken from the reference
object itself so that the stand-alone collision step (pysdm_b200.collision.Geometric, no PySDM
in the process) interpolates in the very same table.  Inside PySDM the B200 backend uses the
table of the reference's own object; this file is data for the stand-alone path only.

    python oracle/gen_gk_table.py      (needs /root/reference or baseline/_ref)
"""
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle", "refshims"))
sys.path.insert(1, next(p for p in ("/root/reference", os.path.join(ROOT, "baseline", "_ref"))
                        if os.path.isdir(os.path.join(p, "PySDM"))))
os.environ.setdefault("CI", "1")
np.NaN = np.nan

from PySDM import Formulae  # noqa: E402
from PySDM.backends import CPU  # noqa: E402
from PySDM.dynamics.terminal_velocity import GunnKinzer1949  # noqa: E402

if __name__ == "__main__":
    gk = GunnKinzer1949(types.SimpleNamespace(backend=CPU(Formulae())))
    out = os.path.join(ROOT, "pysdm_b200", "data", "gunn_kinzer_1949.npz")
    np.savez(out, a=np.asarray(gk.a.data), b=np.asarray(gk.b.data), factor=float(gk.factor),
             maximum_radius=float(gk.maximum_radius))
    print(out, gk.a.data.shape, gk.factor, gk.maximum_radius)
