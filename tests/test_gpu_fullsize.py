"""Hierarchy and solve parity at the HEADLINE size (BASELINE.json configs[1]: 7-pt Laplace 256^3 on one B200): the GPU path
through the C ABI against the CPU oracle on the same 16.8 M-row system, level by level -- sizes, CF splitting, row pointers
and column order of every A_l and P_l bit-for-bit, values within 1e-10 relative (north_star), iterations within +-1, final
relative residual <= the oracle's, solution within 1e-10.  The 27-point reference-generator stencil is checked at 128^3.
Arrays are compared one level and one matrix at a time so that the host never holds two copies of the hierarchy."""
import numpy as np
import pytest

from oracle_lib import Oracle

pytestmark = pytest.mark.gpu

VAL_RTOL = 1e-10


@pytest.mark.parametrize("n,stencil,theta", [(256, 7, 0.25), (128, 27, 0.57)])
def test_hierarchy_bit_exact_fullsize(b200, n, stencil, theta):
    import os
    s = b200.HypreSystem({"linear_system": {"type": f"build_{stencil}pt_stencil", "nx": n, "ny": n, "nz": n},
                          "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50,
                                              "max_iterations": 200, "orthogonalization": "mgs"},
                          "boomeramg_settings": {"strong_threshold": theta}})
    s.setup_precon_and_solver()
    s.load()
    s.solve()
    o = Oracle(tolerance=1e-8, kspace=50, max_iterations=200, strong_threshold=theta)
    o.threads(len(os.sched_getaffinity(0)))
    o.gen_laplace(n, stencil=stencil)
    o.setup()
    xo = o.solve()
    amg = s.precond
    assert amg.num_levels == o.num_levels, (amg.num_levels, o.num_levels)
    for l in range(o.num_levels):
        for which in (("A", "P") if l + 1 < o.num_levels else ("A",)):
            gp, gc, gv = amg.level_matrix(l, which)
            rp, rc, rv, shape = o.matrix(l, which)
            assert len(gp) - 1 == shape[0], f"{which}[{l}]: {len(gp) - 1} rows, oracle {shape[0]}"
            assert np.array_equal(gp, rp), f"{which}[{l}]: row pointers differ"
            assert np.array_equal(gc, rc.astype(gc.dtype)), f"{which}[{l}]: column pattern / entry order differs"
            if len(rv):
                err = float(np.abs(gv - rv).max() / np.abs(rv).max())
                assert err <= VAL_RTOL, f"{which}[{l}]: values differ by {err:.3e} relative"
            del gp, gc, gv, rp, rc, rv
        if l + 1 < o.num_levels:
            assert np.array_equal(amg.level_cf(l), o.cf(l)), f"CF[{l}] differs"
    x = s.solution()
    assert abs(s.solver.iterations - o.iterations) <= 1, (s.solver.iterations, o.iterations)
    assert s.solver.converged == 1
    assert s.solver.relres <= max(o.relres * (1 + 1e-6), 1e-8)
    assert np.abs(x - 1.0).max() < 1e-5            # generator known answer (laplace_3d_weak_scaling.hpp:321)
    if s.solver.iterations == o.iterations:
        assert np.abs(x - xo).max() / np.abs(xo).max() <= VAL_RTOL
    s.destroy_system()
