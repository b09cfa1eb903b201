"""Per-iteration solve time as a function of the number of AMG levels (where does the V-cycle time go, 1 GPU vs N GPUs):
    python profiles/level_sweep.py                       # one GPU
    torchrun --nproc-per-node 2 profiles/level_sweep.py  # one rank per GPU
Prints one JSON line per max_levels value (rank 0)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))


def main():
    import torch
    import b200hypre as bh
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl")
        bh.comm_init_torch()
    else:
        bh.comm_init(0, 1)
    bh.pool_reserve(48 * 2 ** 30)
    n = int(os.environ.get("SIZE", 256))
    stream = torch.cuda.ExternalStream(bh.stream_ptr())
    for ml in (1, 2, 3, 4, 5, 6, 8, 20):
        s = bh.HypreSystem({"linear_system": {"type": "build_7pt_stencil", "nx": n, "ny": n, "nz": n},
                            "solver_settings": {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50, "max_iterations": 40},
                            "boomeramg_settings": {"max_levels": ml}}, rank, world)
        s.setup_precon_and_solver()
        s.load()
        s.assemble_system()
        s.solver.set_precond(s.precond)
        s.solver.setup(s.A, s.b, s.x)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        bh.synchronize()
        e0.record(stream)
        s.solver.solve(s.A, s.b, s.x)
        e1.record(stream)
        bh.synchronize()
        torch.cuda.synchronize()
        if rank == 0:
            it = max(1, s.solver.iterations)
            print(json.dumps({"ranks": world, "max_levels": ml, "levels": s.precond.num_levels, "iterations": s.solver.iterations,
                              "solve_ms": e0.elapsed_time(e1), "per_iteration_ms": e0.elapsed_time(e1) / it}), flush=True)
        s.destroy_system()


if __name__ == "__main__":
    main()
