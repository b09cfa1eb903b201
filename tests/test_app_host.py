"""CPU tests of the C++ host application (hypre-mini-app_b200/app): the YAML-subset parser against PyYAML on
hypre_app.yaml-style inputs, and the no-GPU behaviour of hypre_app (fails loudly, no CPU fallback)."""
import os
import subprocess

import pytest
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
APP = os.path.join(ROOT, "hypre-mini-app_b200", "app")
BIN = os.path.join(ROOT, "hypre-mini-app_b200", "bin", "hypre_app")

SAMPLE = """# -*- mode: yaml -*-
#
# sample input in the layout of etc/hypre_app.yaml
linear_system:
  type: matrix_market         # File type: [matrix_market, hypre_ij]
  matrix_file: "mat # 1.mm"   # quoted, with a hash inside
  rhs_file: rhs.mm
  write_outputs: false
  num_partitions: 2

solver_settings:
  method: gmres
  preconditioner: boomeramg
  tolerance: 1.0e-6
  max_iterations: 100
  kspace: 10
  csv_profile_file: 'timers.csv'

boomeramg_settings:
  coarsen_type: 8
  strong_threshold: 0.57
  non_galerkin_tol: 0.05
  non_galerkin_level_tols:
    levels: [0, 1, 2]
    tolerances: [0.0, 0.01, 0.05]
  extra_list:
    - 3
    - 4
"""


def flatten(node, path=""):
    out = []
    if isinstance(node, dict):
        for k, v in node.items():
            out += flatten(v, f"{path}.{k}" if path else str(k))
    elif isinstance(node, list):
        for i, v in enumerate(node):
            out += flatten(v, f"{path}[{i}]")
    else:
        out.append((path, node))
    return out


@pytest.fixture(scope="module")
def yaml_dump(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("bin") / "yaml_dump")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I", APP, os.path.join(ROOT, "tests", "cpp", "yaml_dump.cpp"), "-o", exe])
    return exe


def test_yaml_subset_matches_pyyaml(yaml_dump, tmp_path):
    f = tmp_path / "in.yaml"
    f.write_text(SAMPLE)
    out = subprocess.run([yaml_dump, str(f), "typed"], capture_output=True, text=True, check=True).stdout.splitlines()
    got = dict(l.split(" = ", 1) for l in out)
    ref = flatten(yaml.safe_load(SAMPLE))
    for path, val in ref:
        assert path in got, path
        if isinstance(val, bool):
            assert got[path] == ("true" if val else "false")
        elif isinstance(val, (int, float)):
            assert float(got[path]) == float(val), path
        else:
            assert got[path] == str(val), path
    assert float(got["tol*2"]) == 2e-6 and got["kspace+1"] == "11" and got["missing"] == "0" and got["flag"] == "0"


@pytest.mark.parametrize("bad", ["a: 1\n   b: 2\n", "key_without_colon\n", "a:\n\tb: 1\n"])
def test_yaml_rejects_malformed(yaml_dump, tmp_path, bad):
    f = tmp_path / "bad.yaml"
    f.write_text(bad)
    r = subprocess.run([yaml_dump, str(f)], capture_output=True, text=True)
    assert r.returncode != 0 and "yaml" in r.stderr


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU behaviour")
def test_app_fails_loudly_without_gpu(tmp_path):
    if not os.path.exists(BIN):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "hypre-mini-app_b200")])
        subprocess.check_call(["make", "-s", "-C", APP])
    f = tmp_path / "in.yaml"
    f.write_text(SAMPLE)
    r = subprocess.run([BIN, str(f)], capture_output=True, text=True)
    assert r.returncode != 0
    assert "no CPU fallback" in r.stderr
