"""Step ne30 L72 variants one dynamics step at a time on the GPU and report where an error latches.  Debug helper."""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from e3sm_b200.params import config, weak_scaled
from e3sm_b200.dycore import ThetaLDycore, build_case
from e3sm_b200 import lib
L = lib.load()
L.thetal_b200_set_abort_on_error(ctypes.byref(ctypes.c_int(0)))
nsteps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
variants = {
    "C2": config("C2"),
    "LIT": config("C2").with_(nu_div=2.5e15),
    "W2 ne42": weak_scaled(config("C2"), 42),
    "W4 ne60": weak_scaled(config("C2"), 60),
    "W8 ne85": weak_scaled(config("C2"), 85),
    "C5 (hvs=3)": config("C5"),
    "C2 nu_top=0": config("C2").with_(nu_top=0.0, hypervis_subcycle_tom=0),
    "C2 nu_div=1e15": config("C2").with_(nu_div=1e15),
    "C2 hvs=2": config("C2").with_(hypervis_subcycle=2),
}
only = sys.argv[2] if len(sys.argv) > 2 else None
for name, p in variants.items():
    if only is not None and not name.startswith(only):
        continue
    grid, hv, st, phis, gradphis, refs = build_case(p)
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
    d.set_state(st)
    nm1, n0, np1 = 0, 1, 2
    print("==", name, flush=True)
    for i in range(nsteps):
        try:
            ms, nm1, n0, np1 = d.time_steps(p.tstep, 1, nm1, n0, np1)
        except RuntimeError as e:
            print("step", i, "ERROR", e, flush=True)
            break
        if i % 50 == 49 or i == nsteps - 1:
            s = d.get_state()
            w = np.abs(s.w_i[:, n0])
            ie, k, pt = np.unravel_index(np.argmax(w), w.shape)
            v = np.abs(s.v[:, n0])
            vi = np.unravel_index(np.argmax(v), v.shape)
            print(f"step {i} ms={ms:.2f} min dp3d={s.dp3d[:, n0].min():.4g} max|w|={w.max():.4g} at (ie={ie},k={k},p={pt}) "
                  f"max|v|={v.max():.4g} at {tuple(int(x) for x in vi)}", flush=True)
