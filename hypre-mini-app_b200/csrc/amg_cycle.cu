// amg_cycle.cu -- BoomerAMG V(1,1)-cycle as GMRES preconditioner, plus BoomerAMG as a stand-alone solver.
// Replaces HYPRE_BoomerAMGSolve (/root/reference/src/HypreSystem.cpp:324; cycle definition SURVEY.md A.7-A.9).
//
// Every smoother sweep is one fused pass over the level operator (SpMV with epilogue, spmv.cu):
//   l1-Jacobi / Jacobi (relax 18 / 7):  u' = u + w (f - A u) ./ d   (out of place, ping-pong buffers; the first
//                                       down sweep starts from u = 0 and degenerates to u = w f ./ d)
//   Chebyshev (relax 16), order o:      r = D^-1/2 (f - A u); v = c_{o-1} r; v = c_k r + D^-1/2 A D^-1/2 v; u += D^-1/2 v
//                                       = 1 residual pass + (o-1) Horner passes, the D^-1/2 scalings folded into the
//                                       gather (xs) and the epilogue.
// Restriction uses the explicitly stored R = P^T (keep_transpose), prolongation adds in place.  Coarsest level:
// the unpivoted Gaussian elimination of SURVEY.md A.9, factored once at setup and applied by one thread block.
// Levels >= 1 (launch-latency bound, internal buffers only) are captured once as a CUDA graph and replayed.
#include "amg.h"

namespace b200 {

namespace {

// x_out[0 .. n_out) = solution entries [out_start, out_start + n_out) (this rank's rows of the replicated coarse solve)
__global__ void k_ge_solve(const double *__restrict__ LU, int n, const double *__restrict__ f, double *__restrict__ x, int out_start,
                           int n_out) {
  extern __shared__ double xs[];
  for (int i = threadIdx.x; i < n; i += blockDim.x) xs[i] = f[i];
  __syncthreads();
  if (n == 1) {
    if (threadIdx.x == 0 && n_out > 0) { double a = LU[0]; x[0] = (a != 0.0) ? xs[0] / a : xs[0]; }
    return;
  }
  for (int k = 0; k < n - 1; ++k) {
    double xk = xs[k];
    for (int j = k + 1 + threadIdx.x; j < n; j += blockDim.x) {
      double fjk = LU[(size_t)j * n + k];
      if (fjk != 0.0) xs[j] -= fjk * xk;
    }
    __syncthreads();
  }
  for (int k = n - 1; k > 0; --k) {
    double akk = LU[(size_t)k * n + k];
    if (akk != 0.0) {
      if (threadIdx.x == 0) xs[k] /= akk;
      __syncthreads();
      double xk = xs[k];
      for (int j = threadIdx.x; j < k; j += blockDim.x) {
        double ujk = LU[(size_t)j * n + k];
        if (ujk != 0.0) xs[j] -= xk * ujk;
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    double a = LU[0];
    if (a != 0.0) xs[0] /= a;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n_out; i += blockDim.x) x[i] = xs[out_start + i];
}

struct LevelBufs {  // ping-pong state of one level during a cycle
  double *cur, *other;
  const double *f;
};

// one relaxation sweep (possibly two C/F half sweeps).  Jacobi-type sweeps flip b.cur / b.other.
void relax_sweep(Amg &amg, AmgLevel &L, LevelBufs &b, int rtype, bool down, bool zero_guess) {
  const AmgOptions &o = amg.opt;
  ParCsr &A = *L.A;
  int n = A.n;
  // A rank may own no rows of a coarse level (n == 0).  It still walks through the sweep: every kernel wrapper below is a no-op for
  // an empty vector, but the halo exchanges inside par_spmv / gs_halo are collective on the host-staged transport of the N-ranks-on-
  // one-GPU tests, and the buffer flips must stay in step with init_bufs.  (An early return here left the other ranks waiting.)
  if (rtype == 16) {
    int ord = (int)L.cheby_coef.size();
    double *r = L.r.p, *v0 = L.t0.p, *v1 = L.t1.p;
    const double *c = L.cheby_coef.data();
    if (zero_guess) {
      vec_pointwise_mul(r, b.f, L.dsm.p, 1.0, n);
    } else {
      SpmvArgs a; a.x = b.cur; a.y = r; a.b = b.f; a.d = L.dsm.p;
      par_spmv(A, SPMV_RES, a);
    }
    if (ord == 1) {  // u += ds .* (c0 r)
      if (zero_guess) vec_pointwise_mul(b.cur, r, L.dsm.p, c[0], n);
      else { vec_pointwise_mul(v0, r, L.dsm.p, c[0], n); vec_axpy(1.0, v0, b.cur, n); }
      return;
    }
    const double *x = r;
    double a1 = c[ord - 1];
    for (int k = ord - 2; k >= 0; --k) {
      SpmvArgs a;
      a.x = x; a.xs = L.dsm.p; a.b = r; a.d = L.dsm.p; a.a0 = c[k]; a.a1 = a1;
      if (k == 0) { a.y = b.cur; a.g = zero_guess ? nullptr : b.cur; a.pt = 1; }
      else { a.y = (x == v0) ? v1 : v0; a.pt = 0; }
      par_spmv(A, SPMV_CHEB, a);
      x = a.y;
      a1 = 1.0;
    }
    return;
  }
  if (is_two_stage_gs(rtype)) {
    // Two-stage Gauss-Seidel (hypre relax 11 forward / 12 backward, one inner iteration; SURVEY.md section 8f.1 asks for a Gauss-Seidel
    // that suits the device): the triangular solve replaced by the first two terms of its Neumann series, all SpMV-class passes,
    //     r = f - A u;  z = w D^-1 r;  u += z;  z = D^-1 (T z);  u -= z      T = strictly lower / upper triangle of the own diag block
    // In place on b.cur (no buffer flip), no C/F ordering.
    const int side = rtype - 11;
    double *r = L.r.p, *z = L.t0.p, *t = L.t1.p;
    const double *res = b.f;
    if (!zero_guess) {
      SpmvArgs a; a.x = b.cur; a.y = r; a.b = b.f;
      par_spmv(A, SPMV_RES, a);
      res = r;
    }
    vec_pointwise_div(z, res, L.dsm.p, o.relax_weight, nullptr, 0, n);
    if (zero_guess) vec_copy(b.cur, z, n);
    else vec_axpy(1.0, z, b.cur, n);
    SpmvArgs ta; ta.x = z; ta.y = t;
    spmv(L.tri[side], nullptr, L.tri_plan[side], SPMV_AX, ta);
    vec_pointwise_div(z, t, L.dsm.p, 1.0, nullptr, 0, n);
    vec_axpy(-1.0, z, b.cur, n);
    return;
  }
  if (is_gs_relax(rtype)) {
    // hybrid Gauss-Seidel (gs.cu): every pass reads b.cur and writes b.other; the halo is gathered once per call
    const bool fwd = (rtype != 4 && rtype != 14), bwd = (rtype == 4 || rtype == 14 || rtype == 6 || rtype == 8);
    const bool l1 = is_l1_gs_relax(rtype);
    const int npass = (o.relax_order == 1) ? 2 : 1;
    for (int pass = 0; pass < npass; ++pass) {
      int pt = 0;
      if (npass == 2) pt = down ? (pass == 0 ? 1 : -1) : (pass == 0 ? -1 : 1);
      const int ps = (pt == 0) ? 0 : (pt == 1 ? 1 : 2);
      if (zero_guess && pass == 0) vec_fill(b.cur, 0.0, n);
      gs_halo(A, b.cur);
      for (int dir = 0; dir < 2; ++dir) {
        if (dir == 0 ? !fwd : !bwd) continue;
        B200_REQUIRE(n == 0 || L.gs[ps][dir].built, "Gauss-Seidel plan missing for this level (internal error)");
        gs_sweep(A, L.gs[ps][dir], l1, L.dsm.p, b.f, b.cur, b.other);
        std::swap(b.cur, b.other);
      }
    }
    return;
  }
  // Jacobi-type
  int npass = (o.relax_order == 1) ? 2 : 1;
  for (int pass = 0; pass < npass; ++pass) {
    int pt = 0;
    if (npass == 2) pt = down ? (pass == 0 ? 1 : -1) : (pass == 0 ? -1 : 1);
    const int *cf = (npass == 2) ? L.cf.p : nullptr;
    if (zero_guess && pass == 0) {
      vec_pointwise_div(b.other, b.f, L.dsm.p, o.relax_weight, cf, pt, n);
    } else {
      SpmvArgs a;
      a.x = b.cur; a.y = b.other; a.b = b.f; a.d = L.dsm.p; a.a0 = o.relax_weight; a.cf = cf; a.pt = pt;
      par_spmv(A, SPMV_JAC, a);
    }
    std::swap(b.cur, b.other);
  }
}

int flips_per_sweep(const AmgOptions &o, int rtype) {
  if (rtype == 16 || is_two_stage_gs(rtype)) return 0;
  const int npass = (o.relax_order == 1) ? 2 : 1;
  if (is_gs_relax(rtype)) return npass * ((rtype == 6 || rtype == 8) ? 2 : 1);
  return npass;
}

void init_bufs(Amg &amg, const double *rhs, double *z, std::vector<LevelBufs> &bufs) {
  const AmgOptions &o = amg.opt;
  int nl = (int)amg.lev.size();
  bufs.resize(nl);
  for (int l = 0; l < nl; ++l) {
    AmgLevel &L = *amg.lev[l];
    bool coarsest = (l == nl - 1);
    double *A0 = (l == 0) ? z : L.u.p, *B0 = L.t0.p;
    if (o.relax_type[1] == 16 || is_two_stage_gs(o.relax_type[1])) B0 = nullptr;  // Chebyshev / two-stage GS work in place (t0/t1 are scratch)
    // number of buffer flips this level will see, so that the final result lands in A0
    int flips = 0;
    if (coarsest) {
      if (amg.ge_n == 0) flips = o.num_sweeps[1] * flips_per_sweep(o, o.relax_type[1]);
    } else {
      flips = o.num_sweeps[1] * flips_per_sweep(o, o.relax_type[1]) + o.num_sweeps[2] * flips_per_sweep(o, o.relax_type[2]);
    }
    LevelBufs &b = bufs[l];
    b.f = (l == 0) ? rhs : L.f.p;
    if (flips % 2 == 0) { b.cur = A0; b.other = B0; }
    else { b.cur = B0; b.other = A0; }
  }
}

void level_down(Amg &amg, int l, std::vector<LevelBufs> &bufs) {
  const AmgOptions &o = amg.opt;
  int nl = (int)amg.lev.size();
  AmgLevel &L = *amg.lev[l];
  LevelBufs &b = bufs[l];
  int n = L.A->n;
  if (l == nl - 1) {
    if (amg.ge_n > 0) {
      const double *f = b.f;
      if (ctx().nranks > 1) {  // gather the coarsest right-hand side: disjoint rows, zero elsewhere, so the sum is exact
        amg.ge_f.zero();
        if (n) vec_copy(amg.ge_f.p + amg.ge_row_start, b.f, n);
        allreduce_sum_dev(amg.ge_f.p, amg.ge_n);
        f = amg.ge_f.p;
      }
      k_ge_solve<<<1, 256, sizeof(double) * amg.ge_n, stream()>>>(amg.ge_lu.p, amg.ge_n, f, b.cur, amg.ge_row_start, n);
      count_launch();
    } else {
      if (o.num_sweeps[1] == 0) vec_fill(b.cur, 0.0, n);
      for (int s = 0; s < o.num_sweeps[1]; ++s) relax_sweep(amg, L, b, o.relax_type[1], true, s == 0);
    }
    return;
  }
  // pre-smoothing from a zero guess
  if (o.num_sweeps[1] == 0) vec_fill(b.cur, 0.0, n);
  for (int s = 0; s < o.num_sweeps[1]; ++s) relax_sweep(amg, L, b, o.relax_type[1], true, s == 0);
  // residual and restriction
  SpmvArgs a; a.x = b.cur; a.y = L.r.p; a.b = b.f;
  par_spmv(*L.A, SPMV_RES, a);
  SpmvArgs ra; ra.x = L.r.p; ra.y = amg.lev[l + 1]->f.p;
  par_spmv(L.R, SPMV_AX, ra);
}

void level_up(Amg &amg, int l, std::vector<LevelBufs> &bufs) {
  const AmgOptions &o = amg.opt;
  AmgLevel &L = *amg.lev[l];
  LevelBufs &b = bufs[l];
  SpmvArgs pa; pa.x = bufs[l + 1].cur; pa.y = b.cur;
  par_spmv(L.P, SPMV_ADD, pa);
  for (int s = 0; s < o.num_sweeps[2]; ++s) relax_sweep(amg, L, b, o.relax_type[2], false, false);
}

// levels 1 .. nl-1 (down and up): all buffers are internal, so this part is captured once as a CUDA graph
void coarse_part(Amg &amg, std::vector<LevelBufs> &bufs) {
  int nl = (int)amg.lev.size();
  for (int l = 1; l < nl; ++l) level_down(amg, l, bufs);
  for (int l = nl - 2; l >= 1; --l) level_up(amg, l, bufs);
}

}  // namespace

// one sweep of the down-smoother of `level` on caller vectors (bench / parity hook: b200hypre_amg_relax)
void relax_level(Amg &amg, int level, double *u, const double *f) {
  B200_REQUIRE(amg.setup_done && level >= 0 && level < (int)amg.lev.size(), "relax_level: bad level or setup missing");
  AmgLevel &L = *amg.lev[level];
  LevelBufs b{u, L.t0.p, f};
  relax_sweep(amg, L, b, amg.opt.relax_type[1], true, false);
  if (b.cur != u) vec_copy(u, b.cur, L.A->n);
}

void amg_vcycle(Amg &amg, const double *rhs, double *z) {
  B200_REQUIRE(amg.setup_done, "BoomerAMG: Solve called before Setup");
  static int use_graph = -1;
  if (use_graph < 0) {
    const char *e = getenv("B200_CUDA_GRAPH");
    use_graph = (e && atoi(e) == 0) ? 0 : 1;
  }
  int nl = (int)amg.lev.size();
  std::vector<LevelBufs> bufs;
  init_bufs(amg, rhs, z, bufs);
  level_down(amg, 0, bufs);
  if (nl > 2) {
    // multi-rank: replayable only when every halo exchange of levels >= 1 runs over peer memory (device-side sequence
    // numbers, parcsr.cu); the coarsest-level all-reduce is an NCCL call, which is capturable
    bool graph_ok = use_graph && !amg.graph_failed;
    if (graph_ok && ctx().nranks > 1) {
      const char *e = getenv("B200_CUDA_GRAPH_DIST");
      if (e && atoi(e) == 0) graph_ok = false;
      for (int l = 1; l < nl && graph_ok; ++l) {
        AmgLevel &L = *amg.lev[l];
        graph_ok = L.A->halo.p2p && (l == nl - 1 || (L.P.halo.p2p && L.R.halo.p2p));
      }
    }
    if (!graph_ok) {
      coarse_part(amg, bufs);
    } else {
      bool replayed = true;
      if (!amg.graph_exec) {
        cudaGraph_t graph = nullptr;
        unsigned long long before = ctx().launches;
        const std::vector<LevelBufs> saved = bufs;
        B200_CUDA(cudaStreamBeginCapture(stream(), cudaStreamCaptureModeThreadLocal));
        bool captured = true;
        try {
          coarse_part(amg, bufs);
        } catch (...) {
          captured = false;
        }
        cudaError_t ce = cudaStreamEndCapture(stream(), &graph);
        if (captured && ce == cudaSuccess && graph) ce = cudaGraphInstantiate(&amg.graph_exec, graph, 0);
        if (graph) cudaGraphDestroy(graph);
        if (!captured || ce != cudaSuccess || !amg.graph_exec) {  // not capturable here: run the levels directly from now on
          cudaGetLastError();
          amg.graph_exec = nullptr;
          amg.graph_failed = true;
          ctx().launches = before;
          bufs = saved;
          coarse_part(amg, bufs);
          replayed = false;
        } else {
          amg.graph_launches = (int)(ctx().launches - before);
          ctx().launches = before;
          B200_CUDA(cudaGraphLaunch(amg.graph_exec, stream()));
          count_launch(amg.graph_launches);
        }
      }
      else {
        B200_CUDA(cudaGraphLaunch(amg.graph_exec, stream()));
        count_launch(amg.graph_launches);
      }
      // the replayed levels leave their result in u (init_bufs parity); mirror the flips the host did not execute
      if (replayed)
        for (int l = 1; l < nl; ++l)
          if (bufs[l].cur != amg.lev[l]->u.p) std::swap(bufs[l].cur, bufs[l].other);
    }
  } else if (nl == 2) {
    level_down(amg, 1, bufs);
  }
  if (nl > 1) level_up(amg, 0, bufs);
  B200_REQUIRE(bufs[0].cur == z, "V-cycle buffer parity error");
}

void amg_solve(Amg &amg, ParCsr &A, const double *b, double *x, int *iters, double *relres) {
  // hypre_BoomerAMGSolve as a solver (method: boomeramg, HypreSystem.cpp:91-117): cycle while ||b - A x|| / ||b|| >= tol
  // and cycles < max_iter; each cycle is x += M^{-1} (b - A x) with the zero-guess V-cycle (the cycle is affine in x).
  int n = A.n;
  i64 ld = ((i64)n + 1) & ~1LL;
  DVec<double> r(ld), z(ld);
  double bnorm = std::sqrt(dot_host(amg.dw, b, b, n));
  int it = 0;
  double rel = 0;
  while (true) {
    SpmvArgs a; a.x = x; a.y = r.p; a.b = b;
    par_spmv(A, SPMV_RES, a);
    double rn = std::sqrt(dot_host(amg.dw, r.p, r.p, n));
    rel = bnorm ? rn / bnorm : rn;
    if (amg.opt.print_level > 1 && ctx().rank == 0) printf("    Cycle %2d   %e    %e\n", it, rn, rel);
    if (!(rel >= amg.opt.tol) || it >= amg.opt.max_iter) break;
    amg_vcycle(amg, r.p, z.p);
    vec_axpy(1.0, z.p, x, n);
    ++it;
  }
  if (iters) *iters = it;
  if (relres) *relres = rel;
}

}  // namespace b200
