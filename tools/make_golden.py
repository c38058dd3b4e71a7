"""Regression fixture of the CPU oracle (NOT a reference-generated vector: the Fortran build cannot run here,
DESIGN.md section 2).  C1 (ne4 L26): three dynamics steps (tstep_type 9, hyperviscosity, sponge) + the PPM
vertical remap; stored: per-field float64 sums, sums of squares, and 64 sampled values per field, plus the
global diagnostics.  tests/test_oracle_invariants.py::test_oracle_matches_regression_fixture and the GPU
parity suite compare against it, so a silent change of the oracle or the kernels shows up.

    python tools/make_golden.py        # rewrites tests/golden/oracle_c1_3steps_remap.json
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def oracle_run():
    from conftest import global_diagnostics
    from e3sm_b200.dycore import build_case
    from e3sm_b200.params import config
    from oracle.oracle import Oracle
    p = config("C1")
    grid, hv, st, phis, gradphis, refs = build_case(p)
    o = Oracle(grid, hv, p, phis, gradphis, refs, st, nthreads=os.cpu_count() or 1)
    nm1, n0, np1 = 0, 1, 2
    for _ in range(p.rsplit):
        o.prim_advance_exp(p.tstep, nm1, n0, np1)
        nm1, n0, np1 = n0, np1, nm1
    assert o.vertical_remap(n0, p.vert_remap_u_alg) == 0 and o.error_flag == 0
    return st, n0, global_diagnostics(st, n0, grid, hv)


def summarise(st, tl, diag):
    rng = np.random.default_rng(2024)
    out = {"time_level": tl, "diagnostics": list(diag), "fields": {}}
    for f in st.FIELDS:
        a = getattr(st, f)[:, tl].ravel()
        idx = np.sort(rng.choice(a.size, 64, replace=False))
        out["fields"][f] = {"sum": float(np.sum(a)), "sumsq": float(np.sum(a * a)),
                            "idx": idx.tolist(), "val": [float(x) for x in a[idx]]}
    return out


if __name__ == "__main__":
    st, tl, diag = oracle_run()
    path = os.path.join(ROOT, "tests", "golden", "oracle_c1_3steps_remap.json")
    with open(path, "w") as fh:
        json.dump(summarise(st, tl, diag), fh, indent=0)
    print("wrote", path)
