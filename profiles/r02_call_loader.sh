#!/bin/bash
# the multi-threaded text loaders: app tests on the GPU, then the read time of a 7 M-entry MatrixMarket file through hypre_app
cd "$GRAFT_REPO_ROOT"
python - <<'PY'
import numpy as np, scipy.sparse as sp, os, time
n = 100
I = sp.eye(n); T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(n, n))
A = (sp.kron(sp.kron(T, I), I) + sp.kron(sp.kron(I, T), I) + sp.kron(sp.kron(I, I), T)).tocoo()
os.makedirs("/tmp/mm", exist_ok=True)
t = time.time()
with open("/tmp/mm/mat.mm", "w") as f:
    f.write("%%MatrixMarket matrix coordinate real general\n" + f"{A.shape[0]} {A.shape[1]} {A.nnz}\n")
    f.write("\n".join(f"{r + 1} {c + 1} {v:.17g}" for r, c, v in zip(A.row.tolist(), A.col.tolist(), A.data.tolist())) + "\n")
b = A @ np.ones(A.shape[0])
with open("/tmp/mm/rhs.mm", "w") as f:
    f.write("%%MatrixMarket matrix array real general\n" + f"{len(b)} 1\n")
    f.write("\n".join(f"{v:.17g}" for v in b.tolist()) + "\n")
open("/tmp/mm/in.yaml", "w").write("linear_system:\n  type: matrix_market\n  matrix_file: mat.mm\n  rhs_file: rhs.mm\nsolver_settings:\n  method: gmres\n  preconditioner: boomeramg\n  tolerance: 1.0e-8\n  max_iterations: 100\n  kspace: 50\n  print_level: 0\nboomeramg_settings:\n  coarsen_type: 8\n  interp_type: 6\n  relax_type: 18\n  relax_order: 0\n  strong_threshold: 0.25\n")
print("files written in", round(time.time() - t, 1), "s,", os.path.getsize("/tmp/mm/mat.mm") / 1e6, "MB,", A.nnz, "entries")
PY
for t in 1 16; do
  (cd /tmp/mm && taskset -c 0-$((t-1)) "$GRAFT_REPO_ROOT"/hypre-mini-app_b200/bin/hypre_app in.yaml 2>&1 | grep -E "read matrix|iterations =|Build HYPRE matrix" | sed "s/^/cores=$t /")
done | tee gpurun_out/r02v_loader_timing.log
