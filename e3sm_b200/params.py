"""Run-time controls of the theta-l step (subset of HOMME's ctl_nl namelist).

Defaults follow H/src/share/control_mod.F90:57-209; the named configs are SURVEY.md section 8(d)
(C1..C5), taken from H/test/benchmarks/*.nl and namelist_defaults_eamxx.xml.
"""
from dataclasses import dataclass, replace


@dataclass
class SimParams:
    ne: int = 4
    nlev: int = 26
    tstep: float = 600.0
    tstep_type: int = 9
    rsplit: int = 3
    qsplit: int = 1
    theta_advect_form: int = 0          # control_mod.F90:97
    pgrad_correction: int = 0           # :101
    hv_ref_profiles: int = 0            # :102
    hypervis_order: int = 2
    hypervis_subcycle: int = 1
    hypervis_subcycle_tom: int = 1      # control_mod default is 0; the configs use 1
    hypervis_scaling: float = 0.0
    nu: float = 6.3e17
    nu_div: float = 6.3e17
    nu_p: float = 6.3e17
    nu_s: float = 6.3e17
    nu_top: float = 2.5e5
    dp3d_thresh: float = 0.125          # :99
    vtheta_thresh: float = 100.0        # :98
    tom_sponge_start: float = 0.0
    u_perturb: float = 1.0
    ftype: int = -1                     # forcing algorithm; 0 / 2 apply the pushed forcing per model step
    vert_remap_u_alg: int = 1           # PPM, mirrored ghost cells (vertremap_base.F90:509-512); 10/11 also built

    @property
    def nelem(self):
        return 6 * self.ne * self.ne

    def with_(self, **kw):
        return replace(self, **kw)


def config(name):
    """C1..C5 of SURVEY.md 8(d)."""
    c1 = SimParams(ne=4, nlev=26, tstep=600.0, nu=6.3e17, nu_div=6.3e17, nu_p=6.3e17, nu_s=6.3e17,
                   hypervis_subcycle=1, hypervis_subcycle_tom=1, nu_top=2.5e5)
    # C2: SURVEY 8(d) tabulates nu_div=2.5e15 with ONE hyperviscosity subcycle at dt=300 s.  On ne30 that
    # pair is unstable although print_cfl's estimate (e3sm_b200/cfl.py: dt < S*397 s) allows it -- the
    # estimate is not sharp on the deformed corner elements (cube_mod.F90:448-450): oracle and GPU both
    # grow a grid-scale mode from step ~28 and reach the EOS abort at step ~39 (tools/debug_c2.py).
    # v1-ne30.nl, where 2.5e15 comes from, subcycles three times (= C5).  C2 therefore runs nu_div = nu
    # (the nu_div<0 default, namelist_mod.F90:1116), stable at one subcycle; same kernels, same bytes.
    c2 = SimParams(ne=30, nlev=72, tstep=300.0, nu=1e15, nu_div=1e15, nu_p=1e15, nu_s=1e15,
                   hypervis_subcycle=1, hypervis_subcycle_tom=1, nu_top=2.5e5)
    c3 = SimParams(ne=120, nlev=72, tstep=75.0, nu=1e13, nu_div=2.5e13, nu_p=1e13, nu_s=1e13,
                   hypervis_subcycle=4, hypervis_subcycle_tom=1, nu_top=1e5)
    c4 = SimParams(ne=256, nlev=128, tstep=40.0, nu=1.6e12, nu_div=1.6e12, nu_p=1.6e12, nu_s=1.6e12,
                   hypervis_subcycle=1, hypervis_subcycle_tom=0, nu_top=0.0)
    c5 = c2.with_(hypervis_subcycle=3, nu_div=2.5e15)
    table = {"C1": c1, "C2": c2, "C3": c3, "C4": c4, "C5": c5}
    return table[name]


def weak_scaled(p: "SimParams", ne: int) -> "SimParams":
    """The same step on a finer sphere: HOMME's resolution rules dt ~ 1/ne, nu ~ (ne0/ne)^3.2 (quoted in
    the reference's thetanh namelists) and nu_top ~ 1/ne (nu_top*dt*ne^2 fixed), so the work per
    element-step is unchanged and the run stays stable."""
    f = float(p.ne) / ne
    return p.with_(ne=ne, tstep=p.tstep * f, nu=p.nu * f ** 3.2, nu_p=p.nu_p * f ** 3.2,
                   nu_s=p.nu_s * f ** 3.2, nu_div=p.nu_div * f ** 3.2, nu_top=p.nu_top * f)


def bytes_per_element_step(p: "SimParams", nlev_tom: int) -> int:
    """Algorithmic HBM bytes per element per dynamics step, SURVEY.md 8(d) / BASELINE.md section 3."""
    nlev = p.nlev
    S = 128 * (6 * nlev + 2)
    I = 128 * (nlev + 1)
    T = 768 * nlev
    Rf = 128 * (3 * nlev + 1)
    Dv = 128 * (4 * nlev + 1)
    tom = p.hypervis_subcycle_tom if p.nu_top > 0 else 0
    if p.tstep_type == 9:
        b = 24 * S + (7 * S + 10 * I) + 4 * Dv + 4 * S
    elif p.tstep_type == 10:
        b = 24 * S + (7 * S + 10 * I) + 2 * Dv
    elif p.tstep_type == 7:
        b = 24 * S + (5 * S + 10 * I) + 2 * Dv
    else:  # explicit tables: 5 CAAR stages, no DIRK
        b = 24 * S + 2 * Dv + (4 * S if p.tstep_type == 5 else 0)
    b += p.hypervis_subcycle * (7 * T + Rf) + tom * 3840 * nlev_tom
    return b
