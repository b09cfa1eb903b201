"""timing helper: 7-point Laplace n^3 with interp_type 0 (classical) or 6 (ext+i): levels, iterations, setup phases, solve time"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))
import numpy as np
import b200hypre as bh
n = int(sys.argv[1]); itype = int(sys.argv[2]) if len(sys.argv) > 2 else 0
bh.comm_init(0, 1)
for rep in range(3):
    s = bh.HypreSystem({"linear_system": {"type": "build_7pt_stencil", "nx": n, "ny": n, "nz": n},
                        "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50, "max_iterations": 400},
                        "boomeramg_settings": {"interp_type": itype, "relax_type": 18, "relax_order": 0, "coarsen_type": 8}})
    s.setup_precon_and_solver(); s.load(); s.solve()
    ph = s.precond.setup_phases() if hasattr(s.precond, "setup_phases") else {}
    print("interp_type", itype, "n", n, "levels", s.precond.num_levels, "its", s.solver.iterations, "relres", s.solver.relres,
          "setup_s", s.precond.setup_seconds, "solve_s", s.solver.solve_seconds, "err", float(np.abs(s.solution() - 1).max()),
          {k: round(v, 2) for k, v in ph.items() if not k.startswith("level.")}, flush=True)
    s.destroy_system()
