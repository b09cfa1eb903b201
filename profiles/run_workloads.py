"""Timings of the other BASELINE.json configurations and of the sibling Krylov methods on one B200 (run under gpurun):
    python profiles/run_workloads.py [--quick] > gpurun_out/workloads_<tag>.jsonl
One JSON line per case: setup / solve seconds (the reference's "Preconditioner setup" and "Solve" timers,
HypreSystem.cpp:685-696,723-734, measured with CUDA events on the library stream), iterations, final relative residual,
max |x - 1| (every system has rhs = A 1) and the SpMV-class kernel bandwidths of the kernel-class profiler.
  lap27_256      : configs[1] with the reference generator's 27-point stencil (theta 0.57), device generator, full size
  poisson27_<n>  : configs[3] perturbed-hex 27-point Poisson + Chebyshev, built on the host (tests/workloads.py) -- reduced
                   size (the 50 M row system needs ~32 GB of host COO and minutes of numpy time; say so next to the numbers)
  momentum3_<n>  : configs[4] 3-DOF nonsymmetric momentum, smoother-only preconditioner (max_levels 1), Jacobi and SGS
  <method>_256   : lap7_256 with cg | bicg | fgmres | cogmres | boomeramg instead of gmres
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def run(bh, name, linear_system, solver, amg, repeat=2):
    import numpy as np
    import torch
    stream = torch.cuda.ExternalStream(bh.stream_ptr())
    best = None
    for k in range(repeat):
        s = bh.HypreSystem({"linear_system": linear_system, "solver_settings": solver, "boomeramg_settings": amg}, *RANKS)
        s.setup_precon_and_solver()
        s.load()
        s.assemble_system()
        bh.profile_enable(True, 8e6)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        ev[0].record(stream)
        s.solver.set_precond(s.precond)
        s.solver.setup(s.A, s.b, s.x)
        ev[1].record(stream)
        s.solver.solve(s.A, s.b, s.x)
        ev[2].record(stream)
        bh.synchronize()
        torch.cuda.synchronize()
        prof = bh.profile_summary()
        bh.profile_enable(False)
        x = s.solution()
        amg_h = s.precond if s.precond is not None else getattr(s, "amg_solver", None)
        rec = {"case": name, "rows": int(s.global_rows), "setup_s": ev[0].elapsed_time(ev[1]) * 1e-3, "solve_s": ev[1].elapsed_time(ev[2]) * 1e-3,
               "iterations": s.solver.iterations, "final_relres": s.solver.relres, "converged": int(s.solver.converged),
               "max_err_vs_1": float(np.abs(x - 1.0).max()), "levels": amg_h.num_levels if amg_h is not None else 0,
               "method": solver["method"], "amg": amg,
               "kernel_gbs": {c: round(v[2] / 1e9 / (v[0] / 1e3)) for c, v in prof.items() if v[0] > 0 and c.startswith(("spmv", "multi", "krylov"))}}
        rec["per_iteration_ms"] = 1e3 * rec["solve_s"] / max(1, rec["iterations"])
        s.destroy_system()
        if best is None or rec["setup_s"] + rec["solve_s"] < best["setup_s"] + best["solve_s"]:
            best = rec
    best["pool_high_water_gib"] = round(bh.pool_stats()["used_high"] / 2 ** 30, 1)
    if RANKS[0] == 0:
        print(json.dumps(best), flush=True)


def main():
    quick = "--quick" in sys.argv
    big = "--big" in sys.argv      # only the two host-built workloads, at the largest size that is practical to build with numpy
    full = "--full" in sys.argv    # BASELINE.json configs[3] / [4] at their STATED sizes, generated on the device
    import b200hypre as bh
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if full:
        return run_full_size(bh, world)
    if "--segregated" in sys.argv:
        k = sys.argv.index("--segregated")
        return run_segregated(bh, int(sys.argv[k + 1]) if len(sys.argv) > k + 1 else 256)
    bh.comm_init(0, 1)
    bh.pool_reserve(int((8 if quick else 60) * 2 ** 30))
    from workloads import momentum3, poisson27_hex
    gm = {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50, "max_iterations": 400}
    n7 = 64 if quick else 256
    if big:
        return run_host_built(bh, gm, 192, 200)
    # lap27 at full size
    run(bh, f"lap27_{n7}", {"type": "build_27pt_stencil", "nx": n7, "ny": n7, "nz": n7}, gm, {"strong_threshold": 0.57})
    # sibling Krylov methods on lap7
    for method in ("gmres", "fgmres", "cogmres", "cg", "bicg", "boomeramg"):
        sv = dict(gm, method=method)
        if method == "boomeramg":
            sv.update(relax_type=18, relax_order=0, strong_threshold=0.25, max_iterations=100)
        run(bh, f"{method}_{n7}", {"type": "build_7pt_stencil", "nx": n7, "ny": n7, "nz": n7}, sv, {})
    run_host_built(bh, gm, 32 if quick else 144, 24 if quick else 128)


def run_full_size(bh, world):
    """configs[3]: perturbed-hex 27-point Poisson, 368^3 = 49.8 M rows, 1.34 G non-zeros, Chebyshev smoother, one GPU.
    configs[4]: 3-DOF momentum, 322^3 x 3 = 100.2 M rows, 2.10 G non-zeros -- beyond 32-bit row pointers on one rank, so it runs
    under torchrun on >= 2 GPUs (block rows per rank); one GPU takes the largest size below 2^31 non-zeros (272^3 x 3 = 60.4 M rows).
        python profiles/run_workloads.py --full
        python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 profiles/run_workloads.py --full"""
    import torch
    rank = int(os.environ.get("RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank))))
        bh.comm_init_torch()
    else:
        bh.comm_init(0, 1)
    gm = {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50, "max_iterations": 400}
    global RANKS
    RANKS = (rank, world)
    if world == 1:
        run(bh, "poisson27_hex_368 (49.8 M rows, device generated)", {"type": "poisson27_hex", "n": 368}, gm,
            {"relax_type": 16, "strong_threshold": 0.5}, repeat=1)
        n = 272
    else:
        n = 322
    tag = f"momentum3_{n} ({3 * n ** 3 / 1e6:.1f} M rows, {world} GPU(s), device generated)"
    run(bh, tag + " gmres+jacobi", {"type": "momentum3", "n": n}, gm, {"max_levels": 1, "relax_type": 18}, repeat=1)
    run(bh, tag + " bicg+jacobi", {"type": "momentum3", "n": n}, dict(gm, method="bicg"), {"max_levels": 1, "relax_type": 18}, repeat=1)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


RANKS = (0, 1)


def run_segregated(bh, n):
    """the reference's own notion of several components (HypreSystem.cpp:1033-1036,681-692; SURVEY.md section 8d input 5): ONE scalar
    matrix, one right-hand side per component, solved one after the other.  The reference calls Setup inside its component loop; this
    times one setup followed by the three solves (what a user of the segregated path pays when the preconditioner is reused)
    and, beside it, the reference's literal order (setup + solve per component).
        python profiles/run_workloads.py --segregated [n]"""
    import numpy as np
    import torch
    bh.comm_init(0, 1)
    stream = torch.cuda.ExternalStream(bh.stream_ptr())
    gm = {"method": "gmres", "preconditioner": "boomeramg", "tolerance": 1e-8, "kspace": 50, "max_iterations": 400}
    s = bh.HypreSystem({"linear_system": {"type": "build_7pt_stencil", "nx": n, "ny": n, "nz": n}, "solver_settings": gm,
                        "boomeramg_settings": {}})
    s.setup_precon_and_solver()
    s.load()
    s.assemble_system()
    N = n ** 3
    g = np.arange(N, dtype=np.float64)
    fields = [np.ones(N), 1.0 + 0.5 * np.sin(2e-5 * g), np.cos(3e-7 * g) + 2.0]     # three "components": known solutions
    xk = bh.IJVector(0, N - 1)
    rhs = []
    for f in fields:
        xk.set_values(f)
        xk.assemble()
        b = bh.IJVector(0, N - 1)
        b.assemble()
        bh.matvec(s.A, xk, b)
        rhs.append(b)
    # untimed warm-up: the first setup grows the device pool and loads the kernels
    s.solver.set_precond(s.precond)
    s.solver.setup(s.A, rhs[0], s.x)
    s.solver.solve(s.A, rhs[0], s.x)
    bh.synchronize()
    for mode in ("one setup, three solves", "setup and solve per component (the reference's loop)"):
        its, errs = [], []
        setup_ms = solve_ms = 0.0
        for c, (f, b) in enumerate(zip(fields, rhs)):
            s.x.fill(0.0)
            if c == 0 or mode.startswith("setup and"):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                s.solver.set_precond(s.precond)
                s.solver.setup(s.A, b, s.x)
                e1.record(stream)
                bh.synchronize(); torch.cuda.synchronize()
                setup_ms += e0.elapsed_time(e1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            s.solver.solve(s.A, b, s.x)
            e1.record(stream)
            bh.synchronize(); torch.cuda.synchronize()
            solve_ms += e0.elapsed_time(e1)
            its.append(int(s.solver.iterations))
            errs.append(float(np.abs(s.solution() - f).max()))
        print(json.dumps({"case": f"segregated_3rhs_lap7_{n}", "mode": mode, "rows": N, "components": 3, "setup_s": setup_ms * 1e-3,
                          "solve_s": solve_ms * 1e-3, "iterations": its, "max_err_vs_known": errs}), flush=True)
    s.destroy_system()


def run_host_built(bh, gm, n_poisson, n_momentum):
    from workloads import momentum3, poisson27_hex
    # perturbed-hex Poisson + Chebyshev (relax 16)
    n = n_poisson
    t0 = time.time()
    A, b = poisson27_hex(n)
    sys.stderr.write(f"poisson27_hex({n}) built in {time.time() - t0:.1f} s\n")
    run(bh, f"poisson27_{n}", {"type": "csr", "ptr": A.indptr, "col": A.indices, "val": A.data, "rhs": b}, gm, {"relax_type": 16, "strong_threshold": 0.5})
    del A, b
    # momentum: smoother-only preconditioner
    n = n_momentum
    t0 = time.time()
    A, b = momentum3(n)
    sys.stderr.write(f"momentum3({n}) built in {time.time() - t0:.1f} s\n")
    ls = {"type": "csr", "ptr": A.indptr, "col": A.indices, "val": A.data, "rhs": b}
    run(bh, f"momentum3_{n}_jacobi", ls, gm, {"max_levels": 1, "relax_type": 18})
    run(bh, f"momentum3_{n}_sgs", ls, gm, {"max_levels": 1, "relax_type": 6, "relax_order": 0})
    run(bh, f"momentum3_{n}_bicg_jacobi", ls, dict(gm, method="bicg"), {"max_levels": 1, "relax_type": 18})


if __name__ == "__main__":
    main()
