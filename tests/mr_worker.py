"""One rank of an N-rank run of the product library (spawned by tests/test_gpu_multirank.py and by hand):
    python tests/mr_worker.py <rank> <nranks> <shm-name|nccl> <out-dir> <json-settings>
All ranks share cuda:0 over the host-staged shm transport (or one GPU each over NCCL when the third argument is "nccl"
and the process runs under torchrun).  Writes <out-dir>/rank<r>.npz with the local pieces of the hierarchy (global
column ids), the solution and the solver statistics."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))


def main():
    rank, nranks, transport, outdir = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4]
    cfg = json.loads(sys.argv[5])
    import b200hypre
    if transport == "nccl":
        rank, nranks = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
        dist.init_process_group("nccl")
        b200hypre.comm_init_torch()
    else:
        b200hypre.comm_init_shm(rank, nranks, transport)
    ls = cfg["linear_system"]
    if ls.get("type") == "csr_npz":   # a global CSR system stored by the caller (profiles/parity_sweep_unstructured.py --ranks)
        z = np.load(ls["file"])
        ls.update(type="csr", ptr=z["ptr"], col=z["col"], val=z["val"], rhs=z["rhs"])
    s = b200hypre.HypreSystem(cfg, rank=rank, nranks=nranks)
    s.setup_precon_and_solver()
    s.load()
    s.solve()
    out = {"x": s.solution(), "row_start": np.int64(s.i_lower), "iterations": np.int64(s.solver.iterations),
           "relres": np.float64(s.solver.relres), "history": np.asarray(s.solver.history())}
    amg = s.precond
    if amg is not None:
        nl = amg.num_levels
        out["num_levels"] = np.int64(nl)
        for l in range(nl):
            kinds = "A" if l + 1 == nl else "APRS"
            for w in kinds:
                ptr, col, val = amg.level_matrix(l, w)
                out[f"{w}{l}_ptr"], out[f"{w}{l}_col"] = ptr, col
                if val is not None:
                    out[f"{w}{l}_val"] = val
            if l + 1 < nl:
                out[f"cf{l}"] = amg.level_cf(l)
            out[f"sm{l}"] = amg.level_smoother(l)
    np.savez(os.path.join(outdir, f"rank{rank}.npz"), **out)
    s.destroy_system()
    b200hypre.lib().b200hypre_comm_finalize()


if __name__ == "__main__":
    main()
