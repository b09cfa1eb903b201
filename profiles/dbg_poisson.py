"""debug helper: setup + solve of the device-generated perturbed-hex Poisson system (or momentum) at size n"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))
import b200hypre as bh
n = int(sys.argv[1]); kind = sys.argv[2] if len(sys.argv) > 2 else "poisson27_hex"
bh.comm_init(0, 1)
amg = {"relax_type": 16, "strong_threshold": 0.5} if kind == "poisson27_hex" else {"max_levels": 1, "relax_type": 18}
s = bh.HypreSystem({"linear_system": {"type": kind, "n": n}, "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8,
                    "kspace": 50, "max_iterations": 400}, "boomeramg_settings": amg})
s.setup_precon_and_solver(); s.load(); s.solve()
import numpy as np
print("n", n, "rows", s.global_rows, "levels", s.precond.num_levels, "its", s.solver.iterations, "relres", s.solver.relres, "err", float(np.abs(s.solution() - 1).max()), flush=True)
s.destroy_system()
