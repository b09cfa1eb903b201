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


def test_reference_arm_ignores_torchrun_thread_cap_and_uses_virtual_ranks():
    """torchrun exports OMP_NUM_THREADS=1: the arm sets its thread count itself (all host cores), and at N > 1 it runs the
    N-virtual-rank oracle on the N-box global problem (round-1 verdict: the N > 1 ratios were void)"""
    cores = len(os.sched_getaffinity(0))
    env = dict(os.environ, RANK="0", LOCAL_RANK="0", WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29572", OMP_NUM_THREADS="1")
    lines = run_reference("--gpus", "2", env=env)
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["n_gpus"] == 2 and d["cpu_baseline"]["cores"] == cores
    assert d["config"]["process_grid"] == [2, 1, 1] and d["config"]["global_rows"] == 2 * 256 ** 3
    assert d["sample_rows"] == 2 * 20 ** 3 and "2 virtual rank(s)" in d["sample"]
    assert d["converged"] is True and d["iterations"] > 0


def test_both_arms_share_one_config():
    sys.path.insert(0, ROOT)
    import bench
    args = bench.parse.__globals__["argparse"].Namespace(n=256, stencil=7, relax_type=18, orth="cgs2")
    c1 = bench.arm_config(args, 1, (1, 1, 1))
    assert c1["rows_per_gpu"] == 256 ** 3 and c1["process_grid"] == [1, 1, 1] and "lap7_256" in c1["workload"]
    # default sample sizes: the GPU arm's own size at N = 1, about the same global size at N > 1
    full = bench.parse.__globals__["argparse"].Namespace(n=256, cpu_n=0, steps=20, warmup=3)
    assert bench.reference_sample(full, 1) == (256, 2, 1)
    assert bench.reference_sample(full, 8)[0] == 128 and bench.reference_sample(full, 2)[0] == 200
