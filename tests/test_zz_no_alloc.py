"""thetal_b200_global_diagnostics_noalloc (include/thetal_b200.h): the device diagnostics with their scratch in a
static device array of the library instead of allocated buffers.  No counterpart in the reference; added
for callers that may not allocate.  Called through the C ABI and compared with
thetal_b200_global_diagnostics bit for bit (both run the same two kernels with an order-independent sum), on two
calls in a row that look at different time levels, so contents left in the shared array by the first call would
show in the second.  Collected last on purpose."""
from ctypes import c_void_p, c_int, byref

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(d, name, tl):
    out = np.full(8, np.nan)
    getattr(d.L, name)(byref(c_int(tl)), c_void_p(out.ctypes.data))
    d._check()
    return out


def test_noalloc_diagnostics_match_the_allocating_entry_point_twice_in_a_row(case_c1):
    from e3sm_b200.dycore import ThetaLDycore
    p, grid, hv, st, phis, gradphis, refs = case_c1
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs)
    try:
        d.set_state(st.copy())
        d.prim_advance_exp(p.tstep, 0, 1, 2)           # time levels 1 and 2 now differ
        want = {tl: _call(d, "thetal_b200_global_diagnostics", tl) for tl in (2, 1)}
        assert want[1][3] != want[2][3]                 # KE of the two levels: the calls below see different data
        first = _call(d, "thetal_b200_global_diagnostics_noalloc", 2)
        second = _call(d, "thetal_b200_global_diagnostics_noalloc", 1)
        third = _call(d, "thetal_b200_global_diagnostics_noalloc", 1)
        for got, tl in ((first, 2), (second, 1), (third, 1)):
            assert np.all(np.isfinite(got))
            assert got.tobytes() == want[tl].tobytes(), (tl, got, want[tl])     # tolerance: none, same bits
        assert first[7] == grid.nelem
        # a bad time level is refused with a message and leaves the output alone
        out = np.full(8, -7.0)
        d.L.thetal_b200_set_abort_on_error(byref(c_int(0)))      # latch the error instead of abort()
        try:
            d.L.thetal_b200_global_diagnostics_noalloc(byref(c_int(99)), c_void_p(out.ctypes.data))
            with pytest.raises(RuntimeError, match="noalloc"):
                d._check()
        finally:
            d.L.thetal_b200_set_abort_on_error(byref(c_int(1)))
        assert np.all(out == -7.0)
        # and the entry point still works after the refusal
        assert _call(d, "thetal_b200_global_diagnostics_noalloc", 2).tobytes() == want[2].tobytes()
    finally:
        d.close()
