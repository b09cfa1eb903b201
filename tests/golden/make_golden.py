#!/usr/bin/env python
"""Generates the golden fixtures of tests/golden/ from the CPU oracle (oracle/oracle.cpp).

The reference's own implementation of this path (hypre) is not vendored in /root/reference and is not installed in
the image, and the reference ships no fixtures, so these vectors cannot come from the reference;
they freeze the behaviour of the pinned oracle (tests/test_oracle_pins.py) so that later edits of the oracle, and the
GPU path compared against it, are checked against committed numbers.  Run from the repository root:
    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import Oracle  # noqa: E402

CASES = {
    "lap7_8": dict(n=8, stencil=7, grid=(1, 1, 1), opts=dict(strong_threshold=0.25, tolerance=1e-8, kspace=50, max_iterations=100)),
    "lap27_6": dict(n=6, stencil=27, grid=(1, 1, 1), opts=dict(strong_threshold=0.57, tolerance=1e-8, kspace=50, max_iterations=100)),
    "lap7_6_2ranks": dict(n=6, stencil=7, grid=(2, 1, 1), opts=dict(strong_threshold=0.25, tolerance=1e-8, kspace=50, max_iterations=100)),
    # classical modified interpolation (interp_type 0, the reference's sample input), on two virtual ranks
    "lap7_8_classical_2ranks": dict(n=(8, 8, 4), stencil=7, grid=(1, 1, 2),
                                    opts=dict(strong_threshold=0.25, interp_type=0, tolerance=1e-8, kspace=50, max_iterations=100)),
}


def main():
    for name, c in CASES.items():
        if len(sys.argv) > 1 and name not in sys.argv[1:]:   # python tests/golden/make_golden.py [case ...]
            continue
        o = Oracle(**c["opts"])
        o.gen_laplace(c["n"], p=c["grid"], stencil=c["stencil"])
        o.setup()
        x = o.solve()
        out = {"n": np.array(c["n"]), "stencil": c["stencil"], "grid": np.array(c["grid"]), "num_levels": o.num_levels,
               "iterations": o.iterations, "x": x, "history": o.history(), "relres": o.relres}
        for k, v in c["opts"].items():
            out["opt_" + k] = v
        for l in range(o.num_levels):
            ptr, col, val, _ = o.matrix(l, "A")
            out[f"A{l}_ptr"], out[f"A{l}_col"], out[f"A{l}_val"] = ptr, col, val
            if l + 1 < o.num_levels:
                out[f"cf{l}"] = o.cf(l)
                ptr, col, val, _ = o.matrix(l, "P")
                out[f"P{l}_ptr"], out[f"P{l}_col"], out[f"P{l}_val"] = ptr, col, val
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, "levels", o.num_levels, "iterations", o.iterations, "relres", o.relres)


if __name__ == "__main__":
    main()
