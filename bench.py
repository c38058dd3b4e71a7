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

    def __init__(self, dev):
        self.dev, self.rows, self.proc = dev, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.dev}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
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
def cpu_reference_sample(p, seconds_target=12.0, warmup=1, steps=None):
    """Time the oracle (reference algorithm restated, -O3, OpenMP over elements) on all host cores
    on a bounded sample of the workload: the same configuration (nlev, stepper, hyperviscosity) on a
    coarser ne8 sphere (384 elements), since the oracle needs a closed sphere for its DSS."""
    from e3sm_b200.dycore import build_case
    from oracle.oracle import Oracle
    cores = os.cpu_count() or 1
    ps = p.with_(ne=8)
    grid, hv, st, phis, gradphis, refs = build_case(ps)
    o = Oracle(grid, hv, ps, phis, gradphis, refs, st, nthreads=cores, fast=True)
    tl = [0, 1, 2]

    def one():
        o.prim_advance_exp(ps.tstep, tl[0], tl[1], tl[2])
        tl[0], tl[1], tl[2] = tl[1], tl[2], tl[0]

    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    one()
    t1 = time.perf_counter() - t0
    n = steps if steps is not None else max(1, min(200, int(seconds_target / max(t1, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(n):
        one()
    el = time.perf_counter() - t0
    assert o.error_flag == 0
    val = grid.nelem * n / el
    sample = (f"ne8np4 L{ps.nlev} sphere (384 elements), {n} steps of the C2 step "
              f"(tstep_type {ps.tstep_type}, hvs {ps.hypervis_subcycle}), {el:.1f} s, oracle -O3 OpenMP")
    return val, cores, sample, el / n


def run_reference(args, world, rank, td):
    if rank != 0:
        return
    p = config(args.config)
    # each bench "step" of this arm is a bounded sample: 20 dynamics steps of the 384-element sphere
    val, cores, sample, sec = cpu_reference_sample(p, warmup=max(1, min(args.warmup, 2)), steps=20 * args.steps)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3 * p.nelem / 384.0,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": f"{args.config}: theta-l NH JW baroclinic wave ne{p.ne}np4 L{p.nlev}, dt={p.tstep:g}s",
                   "note": "reference Fortran+MPI build unavailable on this box (no Fortran compiler); "
                           "CPU restatement of the Fortran routines timed instead; ms_per_step scaled to 5400 elements"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "sypd": p.tstep * (val / p.nelem) / 365.0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args, world, rank, local, td):
    import ctypes
    from e3sm_b200.dycore import ThetaLDycore, broadcast_nccl_id, build_case, _ci, _cd
    from e3sm_b200.initial import State
    from e3sm_b200.partition import Partition
    p = config(args.config)
    if args.weak_ne:
        p = weak_scaled(p, args.weak_ne)      # dt ~ 1/ne, nu ~ (30/ne)^3.2, nu_top ~ 1/ne
    part, nccl_id = None, None
    if world > 1:   # elements cut along HOMME's space-filling curve, one slice per GPU, DSS halos over NCCL
        part = Partition(p.ne, world)
        nccl_id = broadcast_nccl_id(td, rank)
    grid, hv, st, phis, gradphis, refs = build_case(p, part=part, rank=rank)
    d = ThetaLDycore(grid, hv, p, phis, gradphis, refs, device=local, rank=rank, part=part, nccl_id=nccl_id)
    d.set_state(st)
    L = d.L
    nelem, nlev = d.nelem, hv.nlev          # elements on this GPU
    nelem_total = grid.nelem
    K, W = args.steps, max(args.warmup, 3)
    # ---- device-resident timing ----
    ms_w, nm1, n0, np1 = d.time_steps(p.tstep, W, 0, 1, 2)
    l0 = d.launch_count()
    sampler = ClockSampler(local)
    sampler.start()
    barrier(td)
    ms, nm1, n0, np1 = d.time_steps(p.tstep, K, nm1, n0, np1)
    barrier(td)
    clocks = sampler.stop()
    launches = d.launch_count() - l0
    ms = all_max(td, ms)
    value = nelem_total * K / (ms * 1e-3)
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
    alg = {"caar_pre": (14 * S + 4 * Dv) / 5.0, "caar_dss_post": 2 * S, "dirk": (7 * S + 10 * I) / 5.0}
    nl = {"caar_pre": 5, "caar_dss_post": 5, "dirk": 5}
    dom = max(nl, key=lambda k: fam_ms[k])
    dom_ms = fam_ms[dom] / nl[dom]
    ach = nelem * alg[dom] / (dom_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r1_ncu_traffic.json")
    if os.path.exists(tpath) and world == 1:
        tj = json.load(open(tpath))
        if tj.get("workload") == args.config and (dom + "_kernel") in tj:
            traffic = tj[dom + "_kernel"]["traffic_bytes_per_launch"]
    roofline = {"bound": "hbm", "kernel": dom + "_kernel", "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak, "traffic": traffic, "peak_kind": peak_kind,
                "alg_bytes_per_launch": nelem * alg[dom], "launch_ms": dom_ms}
    step_gbs = nelem_total * bstep / (ms / K * 1e-3) / 1e9
    # ---- end to end through the C ABI with host buffers ----
    hst = State(nelem, nlev)
    d.get_state(hst)
    lev = {f: [np.ascontiguousarray(getattr(hst, f)[:, t]) for t in range(3)] for f in State.FIELDS}
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
    e2e_val = nelem_total * Kp / e2e_p
    for f in State.FIELDS:                      # both output buffers hold the step of the same input
        assert np.array_equal(outs[0][f], outs[1][f]) and np.isfinite(outs[0][f]).all(), f
    level_bytes = nelem_total * S
    n_exch, halo_bytes = d.halo_stats()
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, cores, sample, _ = cpu_reference_sample(p)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "weak" if (world == 1 or args.weak_ne) else "strong",
            "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{args.config}{' (weak scaling: ne = 30*sqrt(N), ~5400 elements per GPU)' if args.weak_ne else ''}"
                                   f": theta-l NH JW baroclinic wave ne{p.ne}np4 L{p.nlev}, {nelem_total} elements"
                                   f", tstep_type {p.tstep_type}, dt={p.tstep:g}s, hypervis_subcycle {p.hypervis_subcycle}, sponge on",
                       "l2": "inputs larger than L2 (every kernel streams >= 300 MB of state; L2 is 126 MB)",
                       "parallelism": "1 rank per GPU" + ("" if world == 1 else
                                      f", HOMME space-filling-curve partition into {world} slices ({nelem} elements on rank 0), "
                                      f"DSS halos by ncclSend/ncclRecv: {n_exch} exchanges, {halo_bytes / max(n_exch, 1) / 1e6:.2f} MB "
                                      "sent+received per exchange on rank 0")},
            "sypd": p.tstep * (value / nelem_total) / 365.0,
            "hbm_gbs_step": step_gbs, "hbm_frac_step": step_gbs / (peak * world), "alg_bytes_per_element_step": bstep,
            "kernel_ms_per_step": fam_ms,
            "roofline": roofline,
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": level_bytes, "d2h_bytes_per_step": level_bytes,
                    "steps": Kp, "serial_value": e2e_serial, "serial_steps": Ke,
                    "note": "thetal_b200_pipeline_step: per step pinned-host state(n0) -> device, prim_advance_exp, "
                            "state(np1) -> pinned host; independent batches, the three phases of consecutive steps "
                            "overlap on three streams (timed with the pipeline fill and drain). serial_value: the same "
                            "three phases back to back on one stream"},
            "gpu_launches": launches, "clocks": clocks,
        }
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    d.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default=None,
                    help="C1..C5; default C2. On N>1 GPUs the default is C2 weak-scaled (ne = round(30*sqrt(N)), "
                         "the same work per element and ~5400 elements per GPU, partitioned along the space-filling "
                         "curve); pass --config C3 for the ne120 strong-scaling case")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    world, rank, local, td = dist_setup(args.gpus)
    args.weak_ne = 0
    if args.config is None:
        args.config = "C2"
        if world > 1:
            args.weak_ne = int(round(30.0 * world ** 0.5))      # 2: ne42, 4: ne60, 8: ne85
    if args.impl == "reference":
        if args.steps > 5:
            args.steps = 5          # bounded sample: each step is ~1-4 s of CPU work
        run_reference(args, world, rank, td)
    else:
        run_ours(args, world, rank, local, td)
    if td is not None:
        td.destroy_process_group()


if __name__ == "__main__":
    main()
