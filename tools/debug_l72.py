"""GPU-vs-oracle spot check at L72 (ne4): caar, dirk, full steps.  Debug helper (uses oracle: not product)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from e3sm_b200.params import config
from e3sm_b200.dycore import ThetaLDycore, build_case
from e3sm_b200 import lib
from oracle.oracle import Oracle
import ctypes
L = lib.load()
L.thetal_b200_set_abort_on_error(ctypes.byref(ctypes.c_int(0)))
p = config("C2").with_(ne=4)
grid, hv, st, phis, gradphis, refs = build_case(p)
so = st.copy()
o = Oracle(grid, hv, p, phis, gradphis, refs, so, nthreads=os.cpu_count())
d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
d.set_state(st)
def cmp(tag, tl):
    sg = d.get_state()
    out = []
    for f in ("v", "w_i", "vtheta_dp", "phinh_i", "dp3d"):
        a, b = getattr(sg, f)[:, tl], getattr(so, f)[:, tl]
        den = np.sqrt(np.sum(b * b)) or 1.0
        out.append(f"{f}:{np.sqrt(np.sum((a-b)**2))/den:.2e}")
    print(tag, " ".join(out), "nan" if any(np.isnan(getattr(sg, f)[:, tl]).any() for f in sg.FIELDS) else "", flush=True)
dt = p.tstep
try:
    o.compute_andor_apply_rhs(2, 1, 1, dt / 5, 0.25, 1.0, 0.0, 1.0); d.compute_andor_apply_rhs(2, 1, 1, dt / 5, 0.25, 1.0, 0.0, 1.0); cmp("caar", 2)
    it_o = o.compute_stage_value_dirk(0, 0.0, 1, 0.0, 2, dt / 5); it_g = d.compute_stage_value_dirk(0, 0.0, 1, 0.0, 2, dt / 5); cmp(f"dirk it oracle={it_o} gpu={it_g}", 2)
    o.advance_hypervis(1, dt, 1.0); d.advance_hypervis(1, dt, 1.0); cmp("hypervis", 1)
    a1, a2, a3 = 5 * dt / 18, dt / 36, 8 * dt / 18
    io = o.compute_stage_value_dirk(0, a2, 1, a1, 2, a3); ig = d.compute_stage_value_dirk(0, a2, 1, a1, 2, a3); cmp(f"dirk alpha {io} {ig}", 2)
except RuntimeError as e:
    print("ERROR", e)
