#!/usr/bin/env python
"""Per-stage device time and Newton iteration count of compute_stage_value_dirk inside one tstep_type-9
step (the stage sequence of prim_advance_mod.F90:283-330), on config C2 by default.
  python tools/bench_dirk.py [--config C2] [--reps 5]         (THETAL_B200_LIB / THETAL_B200_DIRK select variants)
"""
import argparse, ctypes, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from e3sm_b200.params import config
from e3sm_b200.dycore import ThetaLDycore, build_case, _ci, _cd

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="C2")
ap.add_argument("--ne", type=int, default=0)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--presteps", type=int, default=2)
a = ap.parse_args()
p = config(a.config)
if a.ne:
    p = p.with_(ne=a.ne)
grid, hv, st, phis, gradphis, refs = build_case(p)
d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
d.set_state(st)
nm1, n0, np1 = 0, 1, 2
for _ in range(a.presteps):
    d.prim_advance_exp(p.tstep, nm1, n0, np1)
    nm1, n0, np1 = n0, np1, nm1
dt, L = p.tstep, d.L
S = 128 * (6 * hv.nlev + 2); I = 128 * (hv.nlev + 1)
def dirk(nm1_, a_nm1, n0_, a_n0, np1_, dt2):
    it = ctypes.c_int(0)
    ms = L.thetal_b200_time_dirk(_ci(nm1_), _cd(a_nm1), _ci(n0_), _cd(a_n0), _ci(np1_), _cd(dt2), _ci(a.reps), ctypes.byref(it))
    d._check()
    return ms, it.value
out = []
caar = d.compute_andor_apply_rhs
caar(nm1, n0, n0, dt / 5, 0.25, 1, 0, 1);        out.append(dirk(nm1, 0.0, n0, 0.0, nm1, dt / 5))
caar(np1, n0, nm1, dt / 5, 0.0, 1, 0, 1);        out.append(dirk(nm1, 0.0, n0, 0.0, np1, dt / 5))
caar(np1, n0, np1, dt / 3, 0.0, 1, 0, 1);        out.append(dirk(nm1, 0.0, n0, 0.0, np1, dt / 3))
caar(np1, n0, np1, 2 * dt / 3, 0.0, 1, 0, 1);    out.append(dirk(nm1, 0.0, n0, 0.0, np1, 2 * dt / 3))
caar(np1, nm1, np1, 3 * dt / 4, 0.75, 1, 0, 1)
out.append(dirk(nm1, dt / 36, n0, 5 * dt / 18, np1, 8 * dt / 18))
tot = sum(m for m, _ in out)
alg = d.nelem * (7 * S + 10 * I)
print(json.dumps({"variant": os.environ.get("THETAL_B200_DIRK", "default"), "lib": os.path.basename(os.environ.get("THETAL_B200_LIB", "default")),
                  "stage_us": [round(m * 1e3, 1) for m, _ in out], "max_iter": [i for _, i in out],
                  "ms_per_step": round(tot, 4), "alg_GBps": round(alg / tot / 1e6, 1)}))
d.close()
