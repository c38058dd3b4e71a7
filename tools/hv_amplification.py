#!/usr/bin/env python
"""Largest amplification factor of one forward-Euler hyperviscosity subcycle on the wind, by power iteration with
advance_hypervis_c on the GPU (show where C2's tabulated nu_div = 2.5e15 stops being stable at one
subcycle).  v <- advance_hypervis(v) = (I - dt nu L) v with L the discrete biharmonic (nu_div/nu on the divergent part);
iterating and renormalising converges to max |1 - dt nu lambda|: > 1 means the subcycle amplifies that mode.
  python tools/hv_amplification.py [--ne 30] [--nlev 26]"""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from e3sm_b200.params import config
from e3sm_b200.dycore import ThetaLDycore, build_case
from e3sm_b200.initial import make_c0

ap = argparse.ArgumentParser()
ap.add_argument("--ne", type=int, default=30)
ap.add_argument("--nlev", type=int, default=26)
ap.add_argument("--iters", type=int, default=60)
a = ap.parse_args()
out = []
for nu_div, hvs in ((1e15, 1), (1.5e15, 1), (2.0e15, 1), (2.5e15, 1), (2.5e15, 3)):
    p = config("C2").with_(ne=a.ne, nlev=a.nlev, nu_div=nu_div, hypervis_subcycle=hvs, nu_top=0.0, hypervis_subcycle_tom=0)
    grid, hv, st, phis, gradphis, refs = build_case(p)
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
    rng = np.random.default_rng(1)
    base = st.copy()
    base.v[:] = 0.0
    v = make_c0(grid, rng.standard_normal(base.v[:, 0].shape))
    amp = 0.0
    for it in range(a.iters):
        s = base.copy()
        s.v[:, 2] = v / np.sqrt((v ** 2).mean())          # unit rms wind perturbation on a resting, reference state
        d.set_state(s)
        d.advance_hypervis(2, p.tstep, 1.0)
        vn = d.get_state().v[:, 2]
        amp = float(np.sqrt((vn ** 2).mean()))
        v = vn
    out.append({"nu_div": nu_div, "hypervis_subcycle": hvs, "dt": p.tstep, "amplification_per_step": amp,
                "stable": amp <= 1.0 + 1e-9})
    d.close()
print(json.dumps({"ne": a.ne, "nlev": a.nlev, "nu": 1e15, "rows": out}, indent=1))
