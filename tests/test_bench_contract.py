"""bench.py's reference arm needs no GPU (it times the CPU restatement on the host cores): run it here and check the
one-line JSON contract the driver parses.  The GPU arm's line is checked on the GPU box by the driver itself."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0"], capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines                        # stdout carries the JSON line and nothing else
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["n_gpus"] == 1 and j["steps"] == 1 and j["warmup"] == 0
    assert j["metric"] == "theta-l dycore element-steps/s" and j["unit"] == "element-steps/s"
    assert j["higher_is_better"] is True and j["dtype"] == "f64" and j["vs_baseline"] is None
    assert "workload" in j["config"] and "ne30np4 L72" in j["config"]["workload"]
    assert j["value"] > 0 and abs(j["value"] - 5400 / (j["ms_per_step"] * 1e-3)) <= 1e-6 * j["value"]
    cb = j["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["sample"] and cb["value"] == j["value"]
    e = j["e2e"]
    assert e["value"] == j["value"] and e["unit"] == j["unit"]
    assert e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


def test_reference_arm_under_torchrun_prints_once_and_every_rank_exits_zero():
    """N > 1: rank 0 alone runs the CPU arm and prints it; the other ranks exit 0 without work (no process group)."""
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "bench.py"),
                          "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip().startswith("{")]
    assert len(lines) == 1, out.stdout[-2000:]
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["n_gpus"] == 2 and j["value"] > 0
