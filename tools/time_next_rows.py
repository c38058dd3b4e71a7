"""Wall-clock (synchronised) cost of the next-row kernels on C2: vertical remap and global diagnostics."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from e3sm_b200.params import config
from e3sm_b200.dycore import ThetaLDycore, build_case
p = config(sys.argv[1] if len(sys.argv) > 1 else "C2")
grid, hv, st, phis, gradphis, refs = build_case(p)
d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
d.set_state(st)
ms, nm1, n0, np1 = d.time_steps(p.tstep, 3, 0, 1, 2)
print(f"3 steps: {ms/3:.3f} ms/step")
for name, fn in (("vertical_remap_c", lambda: d.vertical_remap(n0)), ("global_diagnostics", lambda: d.global_diagnostics(n0))):
    fn()
    t0 = time.perf_counter()
    for _ in range(10):
        fn()
    print(f"{name}: {(time.perf_counter() - t0) / 10 * 1e3:.3f} ms per call (host-synchronised)")
print(d.global_diagnostics(n0))
nstep, f = 0, (1, 2, 3)
d.set_state(st)
t0 = time.perf_counter()
for _ in range(10):
    nstep, *f = d.prim_run_subcycle(p.tstep, nstep, *f)
print(f"prim_run_subcycle_c (rsplit={p.rsplit} steps + remap): {(time.perf_counter() - t0) / 10 * 1e3:.3f} ms per call")
