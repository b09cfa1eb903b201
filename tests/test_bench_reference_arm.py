"""The reference arm of bench.py (`--impl reference`) runs on the host cores only, so its JSON line can be checked here:
same metric / unit / config as the GPU arm, `impl`, a `cpu_baseline` describing the run and an `e2e` object that repeats
the value with zero copy bytes (bench contract).  Uses a small CPU sample so that it takes about a second."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_reference(*extra, env=None):
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1", "--cpu-n", "20", *extra]
    out = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT, timeout=300, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    return lines


def test_reference_arm_prints_one_contract_line():
    lines = run_reference()
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 1
    assert d["unit"] == "Mrow/s" and d["higher_is_better"] is True and d["scaling"] == "weak" and d["dtype"] == "f64"
    assert "workload" in d["config"] and "lap7_256" in d["config"]["workload"]
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and "20^3" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # the sample really solved the system (rhs = A 1)
    assert d["reference_detail"]["relres"] < 1e-8 and d["reference_detail"]["max_err_vs_1"] < 1e-6


def test_reference_arm_runs_on_rank_0_only():
    """under torchrun the driver starts N processes; the other ranks print nothing and exit 0"""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29571")
    assert run_reference("--gpus", "2", env=env) == []
