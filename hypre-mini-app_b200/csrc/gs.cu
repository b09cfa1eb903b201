// gs.cu -- hybrid Gauss-Seidel smoothers of BoomerAMG (relax_type 3, 4, 6, 8, 13, 14) with the exact sequential
// semantics of the CPU algorithm, executed wavefront by wavefront.
// Replaces hypre_BoomerAMGRelax for the Gauss-Seidel family reached from HYPRE_BoomerAMGSolve; relax_type 8 with
// relax_order 1 is the mini-app's DEFAULT smoother (/root/reference/src/HypreSystem.cpp:138,156).  Definition:
// SURVEY.md A.7 -- Gauss-Seidel over a rank's own rows in index order (3 / 13 forward, 4 / 14 backward, 6 / 8 forward
// then backward), off-rank columns frozen at their values before the call (one halo exchange per call), optional C/F
// masking; 3 / 4 / 6: u_i = (f_i - sum_{j != i} a_ij u_j) / a_ii, 8 / 13 / 14: u_i += (f_i - sum_j a_ij u_j) / l1_i with
// hypre's truncated l1 norm  |a_ii| + 1/2 sum_offd |a_ij|  (reset to |a_ii| when <= 4/3 |a_ii|).
//
// B200 formulation.  A sequential sweep is a longest-path problem on the dependency graph "row i needs the new value of
// every updated row j that precedes it and appears in row i"; rows on the same wavefront level are independent.
//   setup  (build_gs_plan): in-degrees from the rows, then Kahn's peeling in ONE persistent kernel (frontier queue in
//          global memory, the transposed pattern gives the dependants, software grid barrier between wavefronts).
//          The queue is the row permutation "sorted by wavefront", the frontier boundaries are the level offsets.
//   sweep  (gs_sweep): ONE persistent kernel per direction; wavefront l = rows order[off[l] .. off[l+1]), G lanes per
//          row, grid barrier between wavefronts.  Values are read from `uout` for rows already updated in this sweep
//          and from `uin` otherwise (out of place: a nonsymmetric pattern cannot create a write-after-read hazard); the
//          row sum (diag block, then offd block) is combined by a shuffle tree, so values agree with the sequential loop
//          to rounding while the choice "updated or not" of every term is exactly the sequential algorithm's.
// The grid never exceeds the number of co-resident blocks, so the software barrier cannot deadlock.
#include "amg.h"
#include <algorithm>

namespace b200 {

void halo_exchange_begin(ParCsr &A, const double *x, const double *xs);
void halo_exchange_end(ParCsr &A);

namespace {

constexpr int GS_THREADS = 256;

__device__ __forceinline__ void grid_barrier(unsigned long long *bar, unsigned nblocks, unsigned long long &epoch) {
  epoch += 1;
  if (nblocks == 1) { __syncthreads(); return; }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1ull);
    const unsigned long long target = epoch * nblocks;
    while (*(volatile unsigned long long *)bar < target) __nanosleep(64);
    __threadfence();
  }
  __syncthreads();
}

__device__ __forceinline__ bool in_set(const int *cf, const double *d, const int *ptr, int i, int points) {
  if (points != 0 && cf[i] != points) return false;
  return ptr[i + 1] > ptr[i] && d[i] != 0.0;
}

// mask[i] = 1 when row i is updated by the sweep; indeg[i] = number of updated rows that precede i (in sweep direction)
// among the columns of row i, -1 for rows outside the set
__global__ void k_gs_indeg(int n, const int *__restrict__ ptr, const int *__restrict__ col, const int *__restrict__ cf,
                           const double *__restrict__ d, int points, int dir, signed char *mask, int *indeg, int *ctr) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  bool member = false;
  if (i < n) {
    member = in_set(cf, d, ptr, i, points);
    int c = -1;
    if (member) {
      c = 0;
      for (int k = ptr[i]; k < ptr[i + 1]; ++k) {
        int j = col[k];
        if ((dir == 0 ? j < i : j > i) && in_set(cf, d, ptr, j, points)) ++c;
      }
    }
    mask[i] = member ? 1 : 0;
    indeg[i] = c;
  }
  int cnt = __syncthreads_count(member);
  if (threadIdx.x == 0 && cnt) atomicAdd(ctr + 5, cnt);   // number of rows in the set
}
// ctr[0] = queue tail (seeds on entry of k_gs_kahn, all scheduled rows on exit), ctr[1..3] = per-level append counters
// (rotating), ctr[4] = depth (output), ctr[5] = rows in the set
__global__ void k_gs_seed(int n, const int *indeg, int *order, int *ctr) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && indeg[i] == 0) order[atomicAdd(ctr, 1)] = i;
}
template <int G>
__global__ void __launch_bounds__(GS_THREADS) k_gs_kahn(const int *__restrict__ Tptr, const int *__restrict__ Tcol, int dir, int *indeg,
                                                        int *order, int *ctr, int *lvl_off, unsigned long long *bar, unsigned nblocks) {
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gthreads = gridDim.x * blockDim.x;
  const int lane = threadIdx.x % G, grp = gtid / G, ngrp = gthreads / G;   // G lanes walk the dependants of one frontier row
  unsigned long long epoch = 0;
  int head = 0, tail = __ldcg(ctr), lvl = 0;
  if (gtid == 0) lvl_off[0] = 0;
  while (head < tail) {
    int *cnt = ctr + 1 + (lvl % 3);
    if (gtid == 0) ctr[1 + ((lvl + 1) % 3)] = 0;   // last read after the barrier of level lvl - 2
    for (int idx = head + grp; idx < tail; idx += ngrp) {
      const int i = __ldcg(order + idx);
      for (int k = Tptr[i] + lane; k < Tptr[i + 1]; k += G) {
        const int j = Tcol[k];
        if (dir == 0 ? j > i : j < i) {
          if (__ldcg(indeg + j) > 0 && atomicSub(indeg + j, 1) == 1) order[tail + atomicAdd(cnt, 1)] = j;
        }
      }
    }
    grid_barrier(bar, nblocks, epoch);
    head = tail;
    tail += __ldcg(cnt);
    ++lvl;
    if (gtid == 0) lvl_off[lvl] = head;
  }
  if (gtid == 0) { ctr[4] = lvl; ctr[0] = tail; }
}

// one Gauss-Seidel pass over the rows of a plan.  G lanes share a row: the entries are loaded and multiplied in parallel
// (one round of dependent loads per row instead of one per entry) and the partial sums are combined by shuffles, so the
// row sum is associated differently from the sequential loop (values agree to rounding, well inside the 1e-10 bar);
// WHICH value every term uses (updated or not) is exactly the sequential algorithm's.
template <int G, bool L1>
__global__ void __launch_bounds__(GS_THREADS) k_gs_sweep(int n, CsrView D, CsrView O, const int *__restrict__ order,
                                                         const int *__restrict__ lvl_off, int depth,
                                                         const signed char *__restrict__ mask, int dir, const double *__restrict__ dsm,
                                                         const double *__restrict__ f, const double *__restrict__ uin, double *uout,
                                                         const double *const *halo_p, unsigned long long *bar, unsigned nblocks) {
  const double *halo = (O.ptr != nullptr) ? *halo_p : nullptr;   // halo values of the exchange that preceded the sweep
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gthreads = gridDim.x * blockDim.x;
  for (int i = gtid; i < n; i += gthreads)
    if (!mask[i]) uout[i] = uin[i];          // rows outside the set keep their value (never read from uout)
  // the row loop is uniform over the warp (its groups take consecutive rows, the tail predicated off), so the reduction
  // shuffles name the full warp; a per-group mask would be serialised over the groups by the compiler
  const int lane = threadIdx.x % G, ngrp = gthreads / G;
  const int wgrp0 = (gtid - (int)(threadIdx.x & 31)) / G, gin = (threadIdx.x & 31) / G;
  unsigned long long epoch = 0;
  for (int l = 0; l < depth; ++l) {
    const int lb = lvl_off[l], le = lvl_off[l + 1];
    for (int idx0 = lb + wgrp0; idx0 < le; idx0 += ngrp) {
      const int idx = idx0 + gin;
      const bool act = idx < le;
      int i = 0, b = 0;
      double sum = 0.0;
      if (act) {
        i = order[idx];
        b = D.ptr[i];
        const int e = D.ptr[i + 1];
        for (int q = (L1 ? b : b + 1) + lane; q < e; q += G) {
          const int c = D.col[q];
          const bool done = mask[c] && (dir == 0 ? c < i : c > i);
          const double v = done ? __ldcg(uout + c) : uin[c];
          sum += D.val[q] * v;
        }
        if (O.ptr != nullptr)
          for (int q = O.ptr[i] + lane; q < O.ptr[i + 1]; q += G) sum += O.val[q] * __ldcg(halo + O.col[q]);
      }
#pragma unroll
      for (int o = G >> 1; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o, G);
      if (act && lane == 0) {
        const double res = f[i] - sum;
        uout[i] = L1 ? uin[i] + res / dsm[i] : res / D.val[b];
      }
    }
    if (l + 1 < depth) grid_barrier(bar, nblocks, epoch);
  }
}

// smoother diagonal of the Gauss-Seidel family: a_ii (3 / 4 / 6) or hypre's truncated l1 norm (8 / 13 / 14)
__global__ void k_gs_diag(int n, CsrView D, CsrView O, int l1, double *d) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b = D.ptr[i], e = D.ptr[i + 1];
  if (b == e) { d[i] = 1.0; return; }
  double dg = D.val[b];
  if (!l1) { d[i] = dg; return; }
  double ad = fabs(dg), s = ad;
  if (O.ptr != nullptr)
    for (int k = O.ptr[i]; k < O.ptr[i + 1]; ++k) s = s + 0.5 * fabs(O.val[k]);
  if (s <= 4.0 / 3.0 * ad) s = ad;
  d[i] = (dg < 0) ? -s : s;
}

int max_resident_blocks(const void *kernel) {
  int per_sm = 0;
  B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, GS_THREADS, 0));
  per_sm = std::max(1, std::min(per_sm, 4));
  return per_sm * ctx().sm_count;
}

}  // namespace

bool is_gs_relax(int t) { return t == 3 || t == 4 || t == 6 || t == 8 || t == 13 || t == 14; }
bool is_l1_gs_relax(int t) { return t == 8 || t == 13 || t == 14; }

void gs_smoother_diag(ParCsr &A, int relax_type, double *d) {
  if (A.n == 0) return;
  CsrView O = A.offd.nnz > 0 ? A.offd.view() : CsrView{0, 0, nullptr, nullptr, nullptr};
  k_gs_diag<<<div_up(A.n, 256), 256, 0, stream()>>>(A.n, A.diag.view(), O, is_l1_gs_relax(relax_type) ? 1 : 0, d);
  count_launch();
}

// T: transposed pattern of A.diag (dependants of a row)
void build_gs_plan(ParCsr &A, const DCsr &T, const int *cf, const double *dsm, int points, int dir, GsPlan &plan) {
  const int n = A.n;
  plan.n = n;
  plan.dir = dir;
  plan.depth = 0;
  plan.width = 0;
  plan.mask.alloc((size_t)std::max(n, 1));
  plan.order.alloc((size_t)std::max(n, 1));
  plan.lvl_off.alloc((size_t)n + 2);
  plan.bar.alloc(2);
  plan.cap = std::min(max_resident_blocks((const void *)k_gs_sweep<32, true>), max_resident_blocks((const void *)k_gs_sweep<8, false>));
  if (n == 0) return;
  DVec<int> indeg((size_t)n), ctr(8);
  ctr.zero();
  plan.bar.zero();
  k_gs_indeg<<<div_up(n, 256), 256, 0, stream()>>>(n, A.diag.ptr.p, A.diag.col.p, cf, dsm, points, dir, plan.mask.p, indeg.p, ctr.p);
  k_gs_seed<<<div_up(n, 256), 256, 0, stream()>>>(n, indeg.p, plan.order.p, ctr.p);
  static int cap = 0;
  if (!cap) cap = std::min(max_resident_blocks((const void *)k_gs_kahn<8>), max_resident_blocks((const void *)k_gs_kahn<32>));
  // frontiers are narrow (rows / depth): a small grid keeps the software barrier cheap
  const double mean = (double)T.nnz / std::max(T.nr, 1);
  const int G = mean <= 10.0 ? 8 : (mean <= 20.0 ? 16 : 32);
  const int nblocks = std::max(1, std::min(std::min(cap, 64), div_up((i64)n * G, (i64)512 * GS_THREADS)));
  if (G == 8)
    k_gs_kahn<8><<<nblocks, GS_THREADS, 0, stream()>>>(T.ptr.p, T.col.p, dir, indeg.p, plan.order.p, ctr.p, plan.lvl_off.p, plan.bar.p, (unsigned)nblocks);
  else if (G == 16)
    k_gs_kahn<16><<<nblocks, GS_THREADS, 0, stream()>>>(T.ptr.p, T.col.p, dir, indeg.p, plan.order.p, ctr.p, plan.lvl_off.p, plan.bar.p, (unsigned)nblocks);
  else
    k_gs_kahn<32><<<nblocks, GS_THREADS, 0, stream()>>>(T.ptr.p, T.col.p, dir, indeg.p, plan.order.p, ctr.p, plan.lvl_off.p, plan.bar.p, (unsigned)nblocks);
  count_launch(3);
  B200_KERNEL_CHECK();
  std::vector<int> h = ctr.to_host();
  plan.depth = h[4];
  plan.rows = h[5];
  B200_REQUIRE(h[0] == h[5], "Gauss-Seidel plan: dependency cycle (internal error)");
  if (plan.depth > 0) {
    std::vector<int> off((size_t)plan.depth + 1);
    B200_CUDA(cudaMemcpyAsync(off.data(), plan.lvl_off.p, sizeof(int) * (plan.depth + 1), cudaMemcpyDeviceToHost, stream()));
    B200_CUDA(cudaStreamSynchronize(stream()));
    for (int l = 0; l < plan.depth; ++l) plan.width = std::max(plan.width, off[l + 1] - off[l]);
    B200_REQUIRE(off[plan.depth] == plan.rows, "Gauss-Seidel plan: level offsets inconsistent (internal error)");
  }
}

template <int G>
static void launch_sweep(ParCsr &A, GsPlan &plan, bool l1, const double *dsm, const double *f, const double *uin, double *uout) {
  const int n = A.n;
  // grid sized for the AVERAGE wavefront (wide ones take a few grid-stride rounds): the cost of a sweep is the number of
  // wavefronts times the barrier latency, which grows with the number of blocks; one block = __syncthreads only
  const int rows_per_block = GS_THREADS / G;
  int want = div_up(n, GS_THREADS);
  if (plan.depth > 0) {
    const int avg = std::max(1, plan.rows / plan.depth);
    want = (avg <= 2 * rows_per_block) ? 1 : div_up(avg, rows_per_block);
  }
  const int nblocks = std::max(1, std::min(plan.cap, want));
  plan.last_blocks = nblocks;
  CsrView O = A.offd.nnz > 0 ? A.offd.view() : CsrView{0, 0, nullptr, nullptr, nullptr};
  B200_CUDA(cudaMemsetAsync(plan.bar.p, 0, sizeof(unsigned long long), stream()));
  if (l1)
    k_gs_sweep<G, true><<<nblocks, GS_THREADS, 0, stream()>>>(n, A.diag.view(), O, plan.order.p, plan.lvl_off.p, plan.depth, plan.mask.p,
                                                              plan.dir, dsm, f, uin, uout, A.xhp(), plan.bar.p, (unsigned)nblocks);
  else
    k_gs_sweep<G, false><<<nblocks, GS_THREADS, 0, stream()>>>(n, A.diag.view(), O, plan.order.p, plan.lvl_off.p, plan.depth, plan.mask.p,
                                                               plan.dir, dsm, f, uin, uout, A.xhp(), plan.bar.p, (unsigned)nblocks);
  count_launch();
}

// no runtime API queries here (plan.cap was taken at plan time): the sweep may be captured into a CUDA graph
void gs_sweep(ParCsr &A, GsPlan &plan, bool l1, const double *dsm, const double *f, const double *uin, double *uout) {
  if (A.n == 0) return;
  const double mean = A.diag.nr > 0 ? (double)A.diag.nnz / A.diag.nr : 1.0;
  if (mean <= 10.0) launch_sweep<8>(A, plan, l1, dsm, f, uin, uout);
  else if (mean <= 20.0) launch_sweep<16>(A, plan, l1, dsm, f, uin, uout);
  else launch_sweep<32>(A, plan, l1, dsm, f, uin, uout);
}

// gathers the halo values of u (the state before the sweep) when the matrix is distributed
void gs_halo(ParCsr &A, const double *u) {
  if (ctx().nranks <= 1) return;
  halo_exchange_begin(A, u, nullptr);
  halo_exchange_end(A);
}

}  // namespace b200
