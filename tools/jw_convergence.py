#!/usr/bin/env python
"""Jablonowski-Williamson (2006) convergence check against the reference-held uncertainty envelope
(tests/golden/jw_reference_uncertainty.json <- H/test/jw_baroclinic/reference_uncertainty.nc): the surface
pressure of two resolutions of this dycore, run on the GPU through prim_run_subcycle_c, must differ in the
l2 norm by less than the spread between the published high-resolution reference models.
  python tools/jw_convergence.py [--ne 30 20] [--days 9] [--nlev 26]
"""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(ne, nlev, days, out_days):
    from e3sm_b200.params import config, weak_scaled
    from e3sm_b200.dycore import ThetaLDycore, build_case
    base = config("C2").with_(nlev=nlev, hypervis_subcycle=3, nu_div=2.5e15)     # v1-ne30.nl settings (C5)
    p = base if ne == 30 else weak_scaled(base, ne)
    grid, hv, st, phis, gradphis, refs = build_case(p)
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
    d.set_state(st)
    per_call = p.rsplit * p.qsplit * p.tstep
    ps = {}
    t, n0 = 0.0, 0
    for day in out_days:
        while t < day * 86400.0 - 1e-6:
            _, _, n0, _ = d.prim_run_subcycle(p.tstep)
            t += per_call
        s = d.get_state()
        ps[day] = hv.hyai[0] * hv.ps0 + s.dp3d[:, n0].sum(axis=1)
    d.close()
    return grid, ps, p


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ne", type=int, nargs=2, default=[30, 20])
    ap.add_argument("--days", type=int, default=9)
    ap.add_argument("--nlev", type=int, default=26)
    a = ap.parse_args()
    from e3sm_b200.interp import evaluate
    env = json.load(open(os.path.join(ROOT, "tests", "golden", "jw_reference_uncertainty.json")))
    days = list(range(1, a.days + 1))
    gA, psA, pA = run(a.ne[0], a.nlev, a.days, days)
    gB, psB, pB = run(a.ne[1], a.nlev, a.days, days)
    rows = []
    for day in days:
        b_on_a = evaluate(gB, psB[day], gA.cart.reshape(-1, 3)).reshape(gA.nelem, 16)
        diff = (psA[day] - b_on_a) / 100.0                                   # hPa
        w = gA.spheremp
        l2 = float(np.sqrt((w * diff ** 2).sum() / w.sum()))
        linf = float(np.abs(diff).max())
        rows.append({"day": day, "l2_hPa": l2, "l2_envelope": env["l2_max"][day], "linf_hPa": linf,
                     "linf_envelope": env["l8_max"][day], "min_ps_hPa": float(psA[day].min() / 100.0)})
    print(json.dumps({"ne": a.ne, "nlev": a.nlev, "dt": [pA.tstep, pB.tstep], "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
