"""Repeat one DIRK call on identical C2 inputs and locate any run-to-run differences.  Debug helper."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from e3sm_b200.params import config
from e3sm_b200.dycore import ThetaLDycore, build_case
p = config("C2")
grid, hv, st, phis, gradphis, refs = build_case(p)
d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
d.set_state(st)
ms, nm1, n0, np1 = d.time_steps(p.tstep, 10, 0, 1, 2)
d.compute_andor_apply_rhs(np1, n0, n0, p.tstep / 5, 0.25, 1.0, 0.0, 1.0)
base = d.get_state()
dt = p.tstep
outs = []
for r in range(8):
    d.set_state(base)
    it = d.compute_stage_value_dirk(nm1, 0.0, n0, 0.0, np1, dt / 5)
    s = d.get_state()
    outs.append((s.w_i[:, np1].copy(), s.phinh_i[:, np1].copy(), it))
for r in range(1, 8):
    dw = outs[r][0] != outs[0][0]
    print("run", r, "iters", outs[r][2], "w diffs", int(dw.sum()), end=" ")
    if dw.any():
        idx = np.argwhere(dw)
        el = np.unique(idx[:, 0]); lv = np.unique(idx[:, 1]); pt = np.unique(idx[:, 2])
        print("elements", el[:8].tolist(), "n_el", len(el), "levels", lv.tolist()[:80], "points", pt.tolist(),
              "max", float(np.abs(outs[r][0] - outs[0][0]).max()))
    else:
        print()
