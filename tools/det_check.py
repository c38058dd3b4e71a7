"""Run the same C2 step several times from the same state: bitwise repeatability + finiteness.  Debug helper."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from e3sm_b200.params import config
from e3sm_b200.dycore import ThetaLDycore, build_case
p = config(sys.argv[1] if len(sys.argv) > 1 else "C2")
if len(sys.argv) > 2:
    p = p.with_(ne=int(sys.argv[2]))
grid, hv, st, phis, gradphis, refs = build_case(p)
d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
d.set_state(st)
ms, nm1, n0, np1 = d.time_steps(p.tstep, int(os.environ.get("DET_STEPS", "40")), 0, 1, 2)
base = d.get_state()
outs = []
for r in range(4):
    d.set_state(base)
    d.prim_advance_exp(p.tstep, nm1, n0, np1)
    s = d.get_state()
    outs.append({f: getattr(s, f)[:, np1].copy() for f in s.FIELDS})
for f in base.FIELDS:
    fin = all(np.isfinite(o[f]).all() for o in outs)
    same = [np.array_equal(outs[0][f], o[f]) for o in outs[1:]]
    mx = max(float(np.abs(outs[0][f] - o[f]).max()) for o in outs[1:])
    print(f, "finite", fin, "repeatable", same, "max diff", mx)
    if not all(same):
        idx = np.argwhere(outs[0][f] != outs[1][f])
        print("   first diffs at", idx[:5].tolist(), "elements", np.unique(idx[:, 0])[:10].tolist(), "count", len(idx))
