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


# ---- multi-threaded body reader of the MatrixMarket / HYPRE-IJ loaders ---------------------------------------------------------
@pytest.fixture(scope="module")
def entries_probe(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("bin") / "mm_entries_probe")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-pthread", "-I", APP, os.path.join(ROOT, "tests", "cpp", "mm_entries_probe.cpp"), "-o", exe])
    return exe


def _reference_parse(text, pattern, cplx, pct):
    """what a single reader with strtoll / strtod makes of the lines (python int / float follow the same grammar for these tokens)"""
    out = []
    for line in text.split("\n"):
        line = line.rstrip("\r")
        if (pct and line.startswith("%")) or not line.strip(" \t"):
            continue
        tok = line.split()
        i, j = int(tok[0]), int(tok[1])
        re = 1.0 if pattern else float(tok[2])
        im = float(tok[3 - (1 if pattern else 0)]) if cplx else 0.0
        out.append((i, j, re, im))
    return out


@pytest.mark.parametrize("eol", ["\n", "\r\n"])
@pytest.mark.parametrize("pattern,cplx", [(0, 0), (0, 1), (1, 0)])
def test_parallel_body_reader_equals_a_single_reader(entries_probe, tmp_path, eol, pattern, cplx):
    """the text loaders cut the body of a file into pieces at line boundaries and parse them on several threads (the reference reads
    with a single fscanf loop); whatever the number of threads, the entries and their ORDER must be those of one reader.  Numbers in
    every spelling the loaders meet: signs, exponents, leading '+', tabs, trailing fields, huge / tiny / out-of-range magnitudes
    (those take the strtod fall-back), comment and blank lines in the middle, last line without terminator."""
    import numpy as np
    rng = np.random.default_rng(5 + pattern + 2 * cplx)
    lines = []
    spell = ["%.17g", "%.3e", "%+.6E", "%.1f", "%g"]
    for k in range(60000):
        i, j = int(rng.integers(1, 10 ** 9)), int(rng.integers(1, 10 ** 9))
        v = float(rng.standard_normal() * 10.0 ** int(rng.integers(-300, 300)))
        w = float(rng.standard_normal())
        f = [str(i) if k % 7 else "+" + str(i), str(j)]
        if not pattern:
            f.append(spell[k % len(spell)] % v)
        if cplx:
            f.append(spell[(k + 2) % len(spell)] % w)
        sep = "\t" if k % 11 == 0 else (" " if k % 3 else "   ")
        line = ("  " if k % 13 == 0 else "") + sep.join(f) + (" trailing 7" if k % 17 == 0 else "") + ("  " if k % 19 == 0 else "")
        lines.append(line)
        if k % 5000 == 1:
            lines += ["% a comment in the middle", "", "   \t "]
    if not pattern:
        lines += ["3 4 1e400" + (" 0" if cplx else ""), "5 6 -1e-400" + (" 0" if cplx else ""), "7 8 0x1p3" + (" 1" if cplx else ""), "9 10 inf" + (" 2" if cplx else "")]
    text = eol.join(lines)      # no terminator after the last line
    path = tmp_path / "body.txt"
    path.write_bytes(text.encode())
    want = _reference_parse(text.replace("0x1p3", "8.0"), pattern, cplx, 1)
    for threads in (1, 2, 3, 8, 32):
        out = subprocess.run([entries_probe, str(path), str(threads), str(pattern), str(cplx), "1"], capture_output=True, text=True, check=True).stdout
        got = [l.split() for l in out.splitlines()]
        assert len(got) == len(want), (threads, len(got), len(want), out[:200])
        gi = np.array([[int(g[0]), int(g[1])] for g in got])
        gv = np.array([[float(g[2]), float(g[3])] for g in got])
        wi = np.array([[w[0], w[1]] for w in want])
        wv = np.array([[w[2], w[3]] for w in want])
        assert np.array_equal(gi, wi), threads
        assert np.array_equal(gv, wv), (threads, np.argwhere(gv != wv)[:3])


def test_parallel_body_reader_reports_malformed_lines(entries_probe, tmp_path):
    body = "\n".join(f"{k + 1} {k + 2} 1.5" for k in range(50000)) + "\n12 oops 3.0\n" + "\n".join(f"{k + 1} {k + 2} 2.5" for k in range(50000)) + "\n"
    path = tmp_path / "bad.txt"
    path.write_text(body)
    for threads in (1, 4):
        out = subprocess.run([entries_probe, str(path), str(threads), "0", "0", "1"], capture_output=True, text=True, check=True).stdout
        assert out.startswith("error Malformed entry line"), out[:100]
    path.write_text("1 2\n")
    out = subprocess.run([entries_probe, str(path), "2", "0", "0", "1"], capture_output=True, text=True, check=True).stdout
    assert out.startswith("error Malformed entry line")
