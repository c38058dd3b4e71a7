#!/usr/bin/env python
"""Benchmark of the theta-l nonhydrostatic dycore step (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            our arm (libthetal_b200.so on B200)
  python bench.py --impl reference --gpus N --steps K ...  reference arm: the reference algorithm's
        CPU restatement (oracle/) on the host cores -- the Fortran+MPI build itself cannot be
        compiled on either box (no Fortran compiler), see DESIGN.md.

A "step" is one prim_advance_exp (5-stage IMEX KGU35 + DIRK + hyperviscosity + sponge) over all
elements of config C2: ne30np4 L72, 5400 elements, JW baroclinic wave, dt=300 s, synthetic state.
Prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from e3sm_b200.params import bytes_per_element_step, config, weak_scaled  # noqa: E402

# stdout carries exactly ONE JSON line: everything else that writes to fd 1 (NCCL's version banner, library
# chatter) is sent to stderr, and the line itself goes to the saved descriptor
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(line):
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


METRIC = "theta-l dycore element-steps/s"
UNIT = "element-steps/s"


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, dev, period_ms=100):
        self.dev, self.rows, self.proc, self.period = dev, [], None, int(period_ms)

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.dev}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", str(self.period)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.th.join(timeout=2)
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for nm, val in zip(names, r[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_setup(ngpus):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    td = None
    if world > 1:
        import torch
        import torch.distributed as td_
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        td_.init_process_group(backend=backend)
        td = td_
    return world, rank, local, td


def all_max(td, x, cuda=True):
    if td is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device="cuda" if (cuda and torch.cuda.is_available()) else "cpu")
    td.all_reduce(t, op=td.ReduceOp.MAX)
    return float(t.item())


def barrier(td):
    if td is not None:
        td.barrier()


# ------------------------------------------------------------------------------------------------
def fortran_toolchain_note():
    """probe for the reference's own toolchain at run time."""
    import shutil
    found = [t for t in ("mpif90", "gfortran", "ifort", "ifx", "nvfortran", "ftn") if shutil.which(t)]
    if found:
        return ("Fortran compiler present (" + ", ".join(found) + ") but the reference build also needs MPI, NetCDF/PIO and "
                "cmake-generated sources: not built; CPU restatement of the Fortran routines timed instead")
    return ("reference Fortran+MPI build unavailable on this box (probed mpif90, gfortran, ifort, ifx, nvfortran, ftn: "
            "none found); CPU restatement of the Fortran routines (oracle/, kind 'port') timed instead")


def cpu_reference_sample(p, seconds_target=12.0, warmup=1, steps=None):
    """Time the oracle (reference algorithm restated in C, -O3, OpenMP over elements) on all host cores on
    the ACTUAL configuration p (same sphere, levels, stepper, hyperviscosity); the sample is bounded in the
    number of dynamics steps, not in the size of the problem."""
    from e3sm_b200.dycore import build_case
    from oracle.oracle import Oracle
    cores = os.cpu_count() or 1
    grid, hv, st, phis, gradphis, refs = build_case(p)
    o = Oracle(grid, hv, p, phis, gradphis, refs, st, nthreads=cores, fast=True)
    tl = [0, 1, 2]

    def one():
        o.prim_advance_exp(p.tstep, tl[0], tl[1], tl[2])
        tl[0], tl[1], tl[2] = tl[1], tl[2], tl[0]

    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    one()
    t1 = time.perf_counter() - t0
    n = steps if steps is not None else max(2, min(200, int(seconds_target / max(t1, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(n):
        one()
    el = time.perf_counter() - t0
    assert o.error_flag == 0
    val = grid.nelem * n / el
    sample = (f"ne{p.ne}np4 L{p.nlev} sphere ({grid.nelem} elements, the configuration itself), {n} dynamics steps "
              f"(tstep_type {p.tstep_type}, hvs {p.hypervis_subcycle}), {el:.1f} s, oracle -O3 OpenMP over elements")
    return val, cores, sample, el / n


def workload_string(name, p, nelem_total, weak_ne=0):
    dev = ""
    if name == "C2":
        dev = (", nu_div = nu = 1e15 (deviates from the SURVEY/BASELINE C2 value nu_div = 2.5e15, which is unstable at "
               "one hyperviscosity subcycle in the oracle and on the GPU alike: DESIGN.md section 6)")
    return (f"{name}{' (weak scaling: ne = 30*sqrt(N), ~5400 elements per GPU)' if weak_ne else ''}"
            f": theta-l NH JW baroclinic wave ne{p.ne}np4 L{p.nlev}, {nelem_total} elements"
            f", tstep_type {p.tstep_type}, dt={p.tstep:g}s, hypervis_subcycle {p.hypervis_subcycle}, "
            f"nu={p.nu:g}, nu_div={p.nu_div:g}, nu_top={p.nu_top:g}, "
            f"sponge {'on' if (p.nu_top > 0 and p.hypervis_subcycle_tom > 0) else 'off'}{dev}")


def run_reference(args, world, rank, td):
    if rank != 0:
        return
    p = config(args.config)
    K = args.steps
    val, cores, sample, sec = cpu_reference_sample(p, warmup=max(1, min(args.warmup, 2)), steps=K)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": K, "warmup": args.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": workload_string(args.config, p, p.nelem), "note": fortran_toolchain_note()},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "sypd": p.tstep * (val / p.nelem) / 365.0,
    }
    emit(line)


def gather_bits(td, vec):
    """all ranks' copies of a small float64 vector, as raw bit patterns (rank-major)."""
    a = np.ascontiguousarray(vec, dtype=np.float64).view(np.int64)
    if td is None:
        return a[None, :]
    import torch
    dev = "cuda" if td.get_backend() == "nccl" else "cpu"
    t = torch.from_numpy(a.copy()).to(dev)
    out = [torch.empty_like(t) for _ in range(td.get_world_size())]
    td.all_gather(out, t)
    return np.stack([o.cpu().numpy() for o in out])


def make_dycore(p, world, rank, local, td):
    from e3sm_b200.dycore import ThetaLDycore, broadcast_nccl_id, build_case
    from e3sm_b200.partition import Partition
    part, nccl_id = None, None
    if world > 1:   # elements cut along HOMME's space-filling curve, one slice per GPU, DSS halos over NCCL
        part = Partition(p.ne, world)
        nccl_id = broadcast_nccl_id(td, rank)
    grid, hv, st, phis, gradphis, refs = build_case(p, part=part, rank=rank)
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs, device=local, rank=rank, part=part, nccl_id=nccl_id)
    d.set_state(st)
    return d, grid, hv


def timed_steps(d, p, td, W, K, tl=(0, 1, 2)):
    """W untimed + K timed dynamics steps with the state resident; device time (CUDA events on the library
    stream), barrier on both sides, max over ranks.  Also checks the answer: dry mass (global mean surface
    pressure, reproducible sum on the device) must not drift, and every rank must hold the same bits."""
    nm1, n0, np1 = tl
    diag0 = d.global_diagnostics(n0)
    _, nm1, n0, np1 = d.time_steps(p.tstep, W, nm1, n0, np1)
    barrier(td)
    ms, nm1, n0, np1 = d.time_steps(p.tstep, K, nm1, n0, np1)
    barrier(td)
    ms = all_max(td, ms)
    diag1 = d.global_diagnostics(n0)
    drift = abs(diag1["ps_mean"] - diag0["ps_mean"]) / diag0["ps_mean"]
    keys = ("ps_min", "ps_max", "ps_mean", "KE", "PE", "IE", "energy")
    bits = gather_bits(td, [diag1[k] for k in keys])
    same = bool((bits == bits[0]).all())
    check = {"dry_mass_rel_drift": drift, "steps": W + K, "ranks_bitwise_identical": same,
             "ps_min": diag1["ps_min"], "energy": diag1["energy"]}
    assert drift <= 1e-11, f"dry mass drifted by {drift:.3e} over {W + K} steps"
    assert same, "global diagnostics differ between ranks"
    assert np.isfinite(diag1["energy"])
    return ms, (nm1, n0, np1), check


def strong_scaling_block(world, rank, local, td, K=5, W=2):
    """BASELINE.json config 3 (ne120np4 L72, hypervis_subcycle 4) on the N GPUs of this run and, on rank 0
    alone, on one GPU: the north-star strong-scaling pair, timed inside the driver-run line."""
    p3 = config("C3")
    d, grid, hv = make_dycore(p3, world, rank, local, td)
    nelem_total = grid.nelem
    ms_n, _, chk = timed_steps(d, p3, td, W, K)
    n_exch, halo_bytes = d.halo_stats()
    d.close()
    ms_1 = 0.0
    if rank == 0:
        d1, _, _ = make_dycore(p3, 1, 0, local, None)
        ms_1, _, _ = timed_steps(d1, p3, None, W, K)
        d1.close()
    ms_1 = all_max(td, ms_1)
    barrier(td)
    return {"config": workload_string("C3", p3, nelem_total), "scaling": "strong", "steps": K, "warmup": W,
            "n1_ms_per_step": ms_1 / K, "nN_ms_per_step": ms_n / K, "n_gpus": world,
            "efficiency": (ms_1 / K) / (world * ms_n / K),
            "value_nN": nelem_total * K / (ms_n * 1e-3), "value_n1": nelem_total * K / (ms_1 * 1e-3),
            "halo_MB_per_exchange_rank0": halo_bytes / max(n_exch, 1) / 1e6, "answer_check": chk}


def run_ours(args, world, rank, local, td):
    import ctypes
    from e3sm_b200.dycore import _ci, _cd
    from e3sm_b200.initial import State
    p = config(args.config)
    if args.weak_ne:
        p = weak_scaled(p, args.weak_ne)      # dt ~ 1/ne, nu ~ (30/ne)^3.2, nu_top ~ 1/ne
    d, grid, hv = make_dycore(p, world, rank, local, td)
    L = d.L
    nelem, nlev = d.nelem, hv.nlev          # elements on this GPU
    nelem_total = grid.nelem
    K, W = args.steps, max(args.warmup, 3)
    # ---- device-resident timing: W warm-up + exactly K timed steps ----
    l0 = d.launch_count()
    sampler = ClockSampler(local, period_ms=20)
    sampler.start()
    ms, (nm1, n0, np1), answer = timed_steps(d, p, td, W, K)
    clocks = sampler.stop()
    launches_per_step = (d.launch_count() - l0 - 4) // (W + K)     # 2 x 2 diagnostics launches
    launches = launches_per_step * K
    value = nelem_total * K / (ms * 1e-3)
    # ---- the same step sustained for >= 1.5 s so the in-line clock record has enough samples ----
    n_sus = max(K, int(1500.0 / max(ms / K, 1e-3)))
    sampler = ClockSampler(local, period_ms=100)
    sampler.start()
    ms_sus, nm1, n0, np1 = d.time_steps(p.tstep, n_sus, nm1, n0, np1)
    clocks_sus = sampler.stop()
    ms_sus = all_max(td, ms_sus)
    clocks["sustained"] = {"steps": n_sus, "ms_per_step": ms_sus / n_sus, "sm_mhz": clocks_sus["sm_mhz"],
                           "reasons": clocks_sus["reasons"], "samples": clocks_sus.get("samples", 0)}
    # ---- per-kernel-family times of one step (CUDA events on the library stream) ----
    fam = (ctypes.c_double * 8)()
    reps = 3
    acc = np.zeros(8)
    for _ in range(reps):
        L.thetal_b200_profile_step(_cd(p.tstep), _ci(nm1), _ci(n0), _ci(np1), fam, ctypes.c_int(8))
        acc += np.array(list(fam))
        nm1, n0, np1 = n0, np1, nm1
    acc /= reps
    names = ["caar_pre", "caar_dss_post", "dirk", "hypervis", "combine", "caar_halo"]
    fam_ms = {n: float(acc[i]) for i, n in enumerate(names)}
    peak, peak_kind = measured_peak()
    S = 128 * (6 * nlev + 2)
    I = 128 * (nlev + 1)
    Dv = 128 * (4 * nlev + 1)
    bstep = bytes_per_element_step(p, d.nlev_tom)
    # algorithmic bytes per launch of each family, tstep_type 9: 5 launches each
    # (the tstep_type 9 combination, 4S, is fused behind the fifth DSS/post launch)
    alg = {"caar_pre": (14 * S + 4 * Dv) / 5.0, "caar_dss_post": (10 * S + (4 * S if p.tstep_type == 9 else 0)) / 5.0,
           "dirk": (7 * S + 10 * I) / 5.0}
    nl = {"caar_pre": 5, "caar_dss_post": 5, "dirk": 5}
    fused_dss = os.environ.get("THETAL_B200_FUSE_DSS", "1") != "0" and os.environ.get("THETAL_B200_DIRK", "1") == "1"
    if fused_dss and p.tstep_type in (9, 10):
        # the DSS of stages 1-4 (2S each: read the pre-DSS buffer, write the state) is done by the DIRK kernel that
        # follows; caar_dss_post_kernel runs once per step (stage 5, with the tstep_type 9 combination behind it)
        alg["caar_dss_post"] = 2 * S + (4 * S if p.tstep_type == 9 else 0)
        alg["dirk"] = (7 * S + 10 * I + 8 * S) / 5.0
        nl["caar_dss_post"] = 1
    kname = {"caar_pre": "caar_pre_kernel", "caar_dss_post": "caar_dss_post_kernel", "dirk": "dirk_col_kernel"}
    dom = max(nl, key=lambda k: fam_ms[k])
    dom_ms = fam_ms[dom] / nl[dom]
    ach = nelem * alg[dom] / (dom_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")
    if os.path.exists(tpath) and world == 1:
        tj = json.load(open(tpath))
        if tj.get("workload") == args.config and kname[dom] in tj:
            traffic = tj[kname[dom]]["traffic_bytes_per_launch"]
    fp64_peak = float(L.thetal_b200_measure_fp64_peak())
    roofline = {"bound": "hbm", "kernel": kname[dom], "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak, "traffic": traffic, "peak_kind": peak_kind,
                "alg_bytes_per_launch": nelem * alg[dom], "launch_ms": dom_ms,
                "note": "average over the family's launches per step (the last DIRK stage also reads state(n0), state(nm1); "
                        "DIRK stages 1-4 include the DSS of the CAAR stage before them); "
                        "the kernel is FP64/XU-pipe bound, not HBM bound: see fp64 and profiles/r2_ncu_summary.txt",
                "by_family": {k: {"launch_ms": fam_ms[k] / nl[k], "GBps": nelem * alg[k] / (fam_ms[k] / nl[k] * 1e-3) / 1e9,
                                  "frac": nelem * alg[k] / (fam_ms[k] / nl[k] * 1e-3) / 1e9 / peak} for k in nl}}
    step_gbs = nelem_total * bstep / (ms / K * 1e-3) / 1e9
    if args.no_e2e:      # capacity runs (C4: 14.5 GB of state per rank): device-resident numbers only
        n_exch, halo_bytes = d.halo_stats()
        d.close()
        if rank == 0:
            emit({"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                  "ms_per_step": ms / K, "higher_is_better": True,
                  "scaling": "weak" if (world == 1 or args.weak_ne) else "strong", "vs_baseline": None,
                  "dtype": "f64", "data": "synthetic",
                  "config": {"workload": workload_string(args.config, p, nelem_total, args.weak_ne),
                             "parallelism": f"1 rank per GPU, {world} ranks, {nelem} elements on rank 0, "
                                            f"{halo_bytes / max(n_exch, 1) / 1e6:.2f} MB per exchange on rank 0"},
                  "sypd": p.tstep * (value / nelem_total) / 365.0, "hbm_gbs_step": step_gbs,
                  "hbm_frac_step": step_gbs / (peak * world), "alg_bytes_per_element_step": bstep,
                  "state_GB_per_gpu": nelem * 3 * S / 1e9, "kernel_ms_per_step": fam_ms, "roofline": roofline,
                  "answer_check": answer, "e2e": None, "gpu_launches": launches, "clocks": clocks})
        return
    # ---- end to end through the reference's own call triple, Fortran-layout host arrays ----
    # init_elements_states_c (H2D, all three time levels) -> prim_run_subcycle_c (rsplit*qsplit dynamics
    # steps + vertical remap, what the Fortran driver calls per model step) -> cxx_push_results_to_f90 (D2H)
    hst = State(nelem, nlev)
    d.get_state(hst)
    for f in State.FIELDS:
        a = getattr(hst, f)
        L.thetal_b200_pin_host(ctypes.c_void_p(a.ctypes.data), ctypes.byref(ctypes.c_long(a.nbytes)), _ci(1))
    d.init_time_level(nm1 + 1, n0 + 1, np1 + 1)
    steps_per_call = p.rsplit * p.qsplit

    def triple():
        d.set_state(hst)
        d.prim_run_subcycle(p.tstep)
        d.get_state(hst)

    triple()
    Kt = max(2, min(6, K))
    barrier(td)
    t0 = time.perf_counter()
    for _ in range(Kt):
        triple()
    e2e_t = all_max(td, time.perf_counter() - t0)
    d._check()
    e2e_val = nelem_total * Kt * steps_per_call / e2e_t
    state_bytes = nelem_total * 3 * S
    assert all(np.isfinite(getattr(hst, f)).all() for f in State.FIELDS)
    for f in State.FIELDS:
        a = getattr(hst, f)
        L.thetal_b200_pin_host(ctypes.c_void_p(a.ctypes.data), ctypes.byref(ctypes.c_long(a.nbytes)), _ci(0))
    # ---- single-level variants (builder-added entry points), kept as labelled extras ----
    nm1, n0, np1 = 0, 1, 2
    lev = {f: [np.ascontiguousarray(getattr(hst, f)[:, t]) for t in range(3)] for f in State.FIELDS}
    del hst
    for f in State.FIELDS:
        for t in range(3):
            a = lev[f][t]
            L.thetal_b200_pin_host(ctypes.c_void_p(a.ctypes.data), ctypes.byref(ctypes.c_long(a.nbytes)), _ci(1))

    def ptr(f, t):
        return ctypes.c_void_p(lev[f][t].ctypes.data)

    def e2e_step(a, b, c):
        L.thetal_b200_push_level(_ci(b), ptr("v", b), ptr("w_i", b), ptr("vtheta_dp", b), ptr("phinh_i", b), ptr("dp3d", b))
        L.thetal_b200_enqueue_step(_cd(p.tstep), _ci(a), _ci(b), _ci(c))
        L.thetal_b200_pull_level(_ci(c), ptr("v", c), ptr("w_i", c), ptr("vtheta_dp", c), ptr("phinh_i", c), ptr("dp3d", c))

    for t in range(3):      # the pulled host levels are consistent with the device again
        L.thetal_b200_push_level(_ci(t), ptr("v", t), ptr("w_i", t), ptr("vtheta_dp", t), ptr("phinh_i", t), ptr("dp3d", t))
    Ke = max(3, min(K, 10))
    for _ in range(2):
        e2e_step(nm1, n0, np1)
        nm1, n0, np1 = n0, np1, nm1
    L.thetal_b200_sync()
    barrier(td)
    t0 = time.perf_counter()
    for _ in range(Ke):
        e2e_step(nm1, n0, np1)
        nm1, n0, np1 = n0, np1, nm1
    L.thetal_b200_sync()
    e2e_s = all_max(td, time.perf_counter() - t0)
    d._check()
    e2e_serial = nelem_total * Ke / e2e_s
    # pipelined service: every step is an independent batch (the pinned host n0 state in, its np1 state out to
    # one of two pinned host buffers); H2D of step j+1, kernels of step j and D2H of step j-1 overlap.
    outs = [{f: np.empty_like(lev[f][0]) for f in State.FIELDS} for _ in range(2)]
    for o_ in outs:
        for a in o_.values():
            L.thetal_b200_pin_host(ctypes.c_void_p(a.ctypes.data), ctypes.byref(ctypes.c_long(a.nbytes)), _ci(1))
    src = n0

    def pipe_step(j):
        o_ = outs[j & 1]
        L.thetal_b200_pipeline_step(_cd(p.tstep), ptr("v", src), ptr("w_i", src), ptr("vtheta_dp", src),
                                    ptr("phinh_i", src), ptr("dp3d", src),
                                    *[ctypes.c_void_p(o_[f].ctypes.data) for f in ("v", "w_i", "vtheta_dp", "phinh_i", "dp3d")])

    Kp = max(K, 10)
    for j in range(4):
        pipe_step(j)
    L.thetal_b200_pipeline_flush()
    barrier(td)
    t0 = time.perf_counter()
    for j in range(Kp):
        pipe_step(j)
    L.thetal_b200_pipeline_flush()
    e2e_p = all_max(td, time.perf_counter() - t0)
    d._check()
    e2e_batch = nelem_total * Kp / e2e_p
    for f in State.FIELDS:                      # both output buffers hold the step of the same input
        assert np.array_equal(outs[0][f], outs[1][f]) and np.isfinite(outs[0][f]).all(), f
    level_bytes = nelem_total * S
    n_exch, halo_bytes = d.halo_stats()
    d.close()
    strong = None
    if world > 1 and args.strong:
        strong = strong_scaling_block(world, rank, local, td)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, cores, sample, _ = cpu_reference_sample(p)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "note": fortran_toolchain_note()}
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "weak" if (world == 1 or args.weak_ne) else "strong",
            "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_string(args.config, p, nelem_total, args.weak_ne),
                       "l2": "inputs larger than L2 (every kernel streams >= 300 MB of state; L2 is 126 MB)",
                       "parallelism": "1 rank per GPU" + ("" if world == 1 else
                                      f", HOMME space-filling-curve partition into {world} slices ({nelem} elements on rank 0), "
                                      f"DSS halos by ncclSend/ncclRecv: {n_exch} exchanges, {halo_bytes / max(n_exch, 1) / 1e6:.2f} MB "
                                      "sent+received per exchange on rank 0")},
            "sypd": p.tstep * (value / nelem_total) / 365.0,
            "hbm_gbs_step": step_gbs, "hbm_frac_step": step_gbs / (peak * world), "alg_bytes_per_element_step": bstep,
            "kernel_ms_per_step": fam_ms,
            "roofline": roofline,
            "fp64": {"peak_tflops_measured": fp64_peak,
                     "how": "independent DFMA chains on every SM (thetal_b200_measure_fp64_peak), CUDA events",
                     "note": "DIRK and CAAR are bound by the FP64 and XU (MUFU/F2F) pipes; pipe utilisation per kernel "
                             "from ncu is in profiles/r2_ncu_summary.txt"},
            "answer_check": answer,
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": state_bytes / steps_per_call,
                    "d2h_bytes_per_step": state_bytes / steps_per_call, "calls": Kt, "dynamics_steps_per_call": steps_per_call,
                    "note": "reference call triple per model step, pinned Fortran-layout host arrays (np,np,[2],nlev,timelevels,nelemd): "
                            "init_elements_states_c (H2D of all three time levels) -> prim_run_subcycle_c (rsplit*qsplit dependent "
                            "dynamics steps + vertical remap) -> cxx_push_results_to_f90 (D2H of all three time levels)",
                    "single_level_serial_value": e2e_serial,
                    "batch_service_value": e2e_batch,
                    "batch_service_note": "thetal_b200_pipeline_step (builder-added): INDEPENDENT inputs, one time level in and out "
                                          f"({level_bytes} B each way), H2D/kernels/D2H of consecutive calls overlapped; not attainable "
                                          "by a time-stepping caller, whose step j+1 consumes step j"},
            "gpu_launches": launches, "clocks": clocks,
        }
        if strong:
            line["strong"] = strong
        if cpu:
            line["cpu_baseline"] = cpu
        emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default=None,
                    help="C1..C5; default C2. On N>1 GPUs the default is C2 weak-scaled (ne = round(30*sqrt(N)), "
                         "the same work per element and ~5400 elements per GPU, partitioned along the space-filling "
                         "curve) as `value`, plus a `strong` block: C3 (ne120np4 L72) on the N GPUs and on one GPU; "
                         "pass --config C3 to make the ne120 strong-scaling case the line's own value")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="device-resident timing only (capacity runs such as C4)")
    ap.add_argument("--no-strong", dest="strong", action="store_false",
                    help="skip the C3 strong-scaling block of the default N>1 run")
    args = ap.parse_args()
    if args.impl == "reference":
        # CPU arm: no process group and no GPU context; under torchrun rank 0 alone runs it, the others exit 0 at once
        if int(os.environ.get("RANK", "0")) != 0:
            return
        if args.config is None:
            args.config = "C2"
        args.steps = min(args.steps, 5)     # bounded sample: each step is ~0.5 s of CPU work on 16 cores
        run_reference(args, int(os.environ.get("WORLD_SIZE", "1")), 0, None)
        return
    world, rank, local, td = dist_setup(args.gpus)
    args.weak_ne = 0
    if args.config is None:
        args.config = "C2"
        if world > 1:
            args.weak_ne = int(round(30.0 * world ** 0.5))      # 2: ne42, 4: ne60, 8: ne85
    else:
        args.strong = False
    run_ours(args, world, rank, local, td)
    if td is not None:
        td.destroy_process_group()


if __name__ == "__main__":
    main()
