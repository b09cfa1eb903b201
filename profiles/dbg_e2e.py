"""debug: does the solution read back through HYPRE_IJVectorGetValues into pinned host memory equal s.solution()?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))
import numpy as np, torch
import b200hypre as bh
bh.comm_init(0, 1)
n = int(sys.argv[1])
cfg = {"linear_system": {"type": "build_7pt_stencil", "nx": n, "ny": n, "nz": n},
       "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50, "max_iterations": 200},
       "boomeramg_settings": {"coarsen_type": 8, "interp_type": 6, "relax_type": 18, "relax_order": 0, "strong_threshold": 0.25}}
s = bh.HypreSystem(cfg); s.setup_precon_and_solver(); s.load(); s.solve()
h_x = torch.full((n ** 3,), -7.0, dtype=torch.float64).pin_memory()
s.x.get_values_ptr(h_x.data_ptr())
bh.synchronize(); torch.cuda.synchronize()
x = s.solution()
print("n", n, "its", s.solver.iterations, "relres", s.solver.relres, "err(solution())", float(np.abs(x - 1).max()), "err(pinned)", float((h_x - 1.0).abs().max()),
      "pinned==solution", bool(np.array_equal(h_x.numpy(), x)), "min/max", float(h_x.min()), float(h_x.max()))
