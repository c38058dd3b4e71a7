#!/usr/bin/env python
"""tests/golden/jw_reference_uncertainty.json from the one numeric file the reference holds for this path:
H/test/jw_baroclinic/reference_uncertainty.nc (Jablonowski & Williamson 2006: l1/l2/linf spread of the surface
pressure, in hPa, between the four high-resolution reference models, days 0..30).  Run in the build container,
where /root/reference exists; the GPU box only sees the JSON."""
import json, os, sys
from scipy.io import netcdf_file
src = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/components/homme/test/jw_baroclinic/reference_uncertainty.nc"
f = netcdf_file(src, "r", mmap=False)
out = {"source": "components/homme/test/jw_baroclinic/reference_uncertainty.nc", "units": "hPa (mb), surface pressure",
       "minmax": f.minmax.decode(), "days": [float(x) for x in f.variables["time"][:]]}
for k in ("l1_max", "l2_max", "l8_max"):
    out[k] = [float(x) for x in f.variables[k][:]]
dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "jw_reference_uncertainty.json")
json.dump(out, open(dst, "w"), indent=1)
print("wrote", dst)
