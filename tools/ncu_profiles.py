#!/usr/bin/env python
"""profiles/<tag>_ncu_summary.txt and profiles/<tag>_ncu_traffic.json from one `ncu --set full` report of a bench
step (run here, no GPU needed):  python tools/ncu_profiles.py gpurun_out/r2_step_c.ncu-rep r2 "command line"  """
import csv, io, json, os, subprocess, sys, collections
rep, tag = sys.argv[1], sys.argv[2]
cmd = sys.argv[3] if len(sys.argv) > 3 else ""
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
cols = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64%"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "warp-inst")]
def tobytes(v, u):
    f = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    return float(v) * f
out = [f"# {tag}: ncu --set full --clock-control none --import-source on, {cmd}",
       "# one dynamics step of C2 (ne30np4 L72, 5400 elements); per launch, cold L2, serialised (compare shares, not absolutes)",
       "# kernel | us | DRAM read MB | DRAM write MB | DRAM % | FP64 pipe % | XU pipe % | issue % | warps active % | regs | warp instructions"]
traffic = collections.OrderedDict()
tot = 0.0
for r in data:
    name = r[ix["Kernel Name"]]
    short = name.replace("void ", "").split("(")[0]
    rd = tobytes(r[ix["dram__bytes_read.sum"]], units[ix["dram__bytes_read.sum"]])
    wr = tobytes(r[ix["dram__bytes_write.sum"]], units[ix["dram__bytes_write.sum"]])
    us = float(r[ix["gpu__time_duration.sum"]])
    if units[ix["gpu__time_duration.sum"]] == "ms": us *= 1e3
    if units[ix["gpu__time_duration.sum"]] == "ns": us /= 1e3
    tot += us
    vals = [f"{us:8.1f}", f"{rd/1e6:7.1f}", f"{wr/1e6:7.1f}"] + [f"{float(r[ix[c]]):6.1f}" for c, _ in cols[3:8]] + \
           [r[ix["launch__registers_per_thread"]], f"{float(r[ix['smsp__inst_executed.sum']])/1e6:7.1f}M"]
    out.append(f"{short[:44]:44s} | " + " | ".join(vals))
    key = short.split("<")[0].replace("thetal::", "")
    t = traffic.setdefault(key, {"launches": 0, "bytes": 0.0, "us": 0.0})
    t["launches"] += 1; t["bytes"] += rd + wr; t["us"] += us
out.append(f"# sum of the {len(data)} launches: {tot:.0f} us")
tj = {"workload": "C2", "source": os.path.basename(rep), "command": cmd}
for k, t in traffic.items():
    tj[k] = {"traffic_bytes_per_launch": t["bytes"] / t["launches"], "launches_per_step": t["launches"],
             "us_per_launch": t["us"] / t["launches"]}
open(os.path.join(ROOT, "profiles", f"{tag}_ncu_summary.txt"), "w").write("\n".join(out) + "\n")
json.dump(tj, open(os.path.join(ROOT, "profiles", f"{tag}_ncu_traffic.json"), "w"), indent=1)
print("\n".join(out[-8:]))
