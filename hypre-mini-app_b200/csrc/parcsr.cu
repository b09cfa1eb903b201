// parcsr.cu -- ParCSR operator: row partition, halo (comm-package) construction and the distributed SpMV-class
// product  y = epilogue(A_diag x_local + A_offd x_halo).
// Replaces hypre's ParCSR comm package + hypre_ParCSRMatrixMatvec reached from HYPRE_ParCSRGMRESSolve /
// HYPRE_BoomerAMGSolve (/root/reference/src/HypreSystem.cpp:723, :324).
//
// Multi-rank product: (1) pack x[send_idx] (scaled by xs when the gather is scaled) on the compute stream,
// (2) grouped NCCL send/recv of the halo segments on the comm stream, overlapped with (3) the diag product with its
// fused epilogue on the compute stream, (4) offd correction on the rows that own offd entries (every epilogue is
// affine in t = sum_j a_ij x_j, so the halo part is added afterwards without a second pass over the matrix).
#include "amg.h"
#include <algorithm>

namespace b200 {

void allgather_host_i64(const i64 *in, i64 *out, int count);
void exchange_dev(int nsend, const int *send_rank, const void *const *send_ptr, const size_t *send_bytes, int nrecv,
                  const int *recv_rank, void *const *recv_ptr, const size_t *recv_bytes, cudaStream_t s);

// ---- pure host logic (also exercised by the CPU/gloo tests through b200hypre_test_* hooks) ---------------------------
// block-row partition of /root/reference/src/HypreSystem.cpp:525-544
void block_partition(i64 n, int np, i64 *starts /* np + 1 */) {
  i64 q = n / np, r = n % np;
  for (int k = 0; k <= np; ++k) starts[k] = q * k + std::min<i64>(k, r);
}
// receive side of the comm package: the sorted halo column ids grouped by owning rank (contiguous segments)
void plan_recv(const i64 *col_map, int n_halo, const i64 *col_starts, int nranks, std::vector<int> &recv_rank,
               std::vector<int> &recv_off) {
  recv_rank.clear();
  recv_off.clear();
  int k = 0;
  for (int r = 0; r < nranks && k < n_halo; ++r) {
    int b = k;
    while (k < n_halo && col_map[k] < col_starts[r + 1]) ++k;
    if (k > b) { recv_rank.push_back(r); recv_off.push_back(b); }
  }
  recv_off.push_back(n_halo);
}

namespace {

__global__ void k_pack(int n, const int *__restrict__ idx, const double *__restrict__ x, const double *__restrict__ xs,
                       double *__restrict__ out) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  int i = idx[k];
  double v = x[i];
  if (xs) v *= xs[i];
  out[k] = v;
}
__global__ void k_global_to_local(int n, const i64 *__restrict__ g, i64 start, int *__restrict__ out) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) out[k] = (int)(g[k] - start);
}

template <int MODE>
__global__ void k_offd_correct(int nrows, const int *__restrict__ rows, CsrView O, const double *__restrict__ xh, SpmvArgs a) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nrows) return;
  int i = rows[t];
  if (MODE == SPMV_JAC && a.cf != nullptr && a.cf[i] != a.pt) return;
  double s = 0.0;
  for (int k = O.ptr[i]; k < O.ptr[i + 1]; ++k) s += O.val[k] * xh[O.col[k]];
  if (MODE == SPMV_AX || MODE == SPMV_ADD) a.y[i] += s;
  else if (MODE == SPMV_RES) a.y[i] -= (a.d ? a.d[i] : 1.0) * s;
  else if (MODE == SPMV_JAC) a.y[i] -= a.a0 * s / a.d[i];
  else { double di = a.d[i]; a.y[i] += (a.pt ? di * di : di) * a.a1 * s; }
}

cudaEvent_t ev_packed = nullptr, ev_halo = nullptr;

}  // namespace

void ParCsr::finalize_spmv() {
  if (!sell.valid()) maybe_build_sell(diag, sell);
  plan = make_plan(diag, &sell);
}

// Fills row_starts / col_starts and, for nranks > 1, the halo plan.  A.col_map_offd, A.row_start, A.col_start, A.n and
// A.n_cols_local must be set.  Collective.
void build_partition_and_halo(ParCsr &A) {
  Context &c = ctx();
  const int np = c.nranks;
  A.row_starts.assign(np + 1, 0);
  A.col_starts.assign(np + 1, 0);
  if (np == 1) {
    A.row_starts[0] = A.row_start; A.row_starts[1] = A.row_start + A.n;
    A.col_starts[0] = A.col_start; A.col_starts[1] = A.col_start + A.n_cols_local;
    return;
  }
  i64 mine[2] = {A.row_start, A.col_start};
  std::vector<i64> all(2 * np);
  allgather_host_i64(mine, all.data(), 2);
  for (int r = 0; r < np; ++r) { A.row_starts[r] = all[2 * r]; A.col_starts[r] = all[2 * r + 1]; }
  A.row_starts[np] = A.g_rows + A.row_starts[0];
  A.col_starts[np] = A.g_cols + A.col_starts[0];
  HaloPlan &h = A.halo;
  plan_recv(A.col_map_offd.data(), A.n_halo, A.col_starts.data(), np, h.recv_rank, h.recv_off);
  // how many entries every rank requests from every other rank
  std::vector<i64> req(np, 0), allreq((size_t)np * np);
  for (size_t k = 0; k < h.recv_rank.size(); ++k) req[h.recv_rank[k]] = h.recv_off[k + 1] - h.recv_off[k];
  allgather_host_i64(req.data(), allreq.data(), np);
  h.send_rank.clear();
  h.send_off.assign(1, 0);
  for (int r = 0; r < np; ++r) {
    i64 cnt = allreq[(size_t)r * np + c.rank];
    if (cnt > 0) { h.send_rank.push_back(r); h.send_off.push_back(h.send_off.back() + (int)cnt); }
  }
  h.n_send = h.send_off.back();
  // ship the requested global ids to their owners
  DVec<i64> d_map, d_req((size_t)std::max(h.n_send, 1));
  d_map.from_host(A.col_map_offd);
  std::vector<const void *> sp;
  std::vector<void *> rp;
  std::vector<size_t> sb, rb;
  for (size_t k = 0; k < h.recv_rank.size(); ++k) {
    sp.push_back(d_map.p + h.recv_off[k]);
    sb.push_back((size_t)(h.recv_off[k + 1] - h.recv_off[k]) * sizeof(i64));
  }
  for (size_t k = 0; k < h.send_rank.size(); ++k) {
    rp.push_back(d_req.p + h.send_off[k]);
    rb.push_back((size_t)(h.send_off[k + 1] - h.send_off[k]) * sizeof(i64));
  }
  exchange_dev((int)h.recv_rank.size(), h.recv_rank.data(), sp.data(), sb.data(), (int)h.send_rank.size(), h.send_rank.data(),
               rp.data(), rb.data(), stream());
  h.send_idx.alloc((size_t)std::max(h.n_send, 1));
  h.send_buf.alloc((size_t)std::max(h.n_send, 1));
  if (h.n_send) {
    k_global_to_local<<<div_up(h.n_send, 256), 256, 0, stream()>>>(h.n_send, d_req.p, A.col_start, h.send_idx.p);
    count_launch();
  }
  A.x_halo.alloc((size_t)std::max(A.n_halo, 1));
  A.x_halo.zero();
  B200_CUDA(cudaStreamSynchronize(stream()));
}

// forward halo exchange of an arbitrary local vector (8-byte elements): halo_out[j] = x_owner[col_map_offd[j]]
void halo_exchange_begin(ParCsr &A, const double *x, const double *xs) {
  Context &c = ctx();
  HaloPlan &h = A.halo;
  if (!ev_packed) {
    B200_CUDA(cudaEventCreateWithFlags(&ev_packed, cudaEventDisableTiming));
    B200_CUDA(cudaEventCreateWithFlags(&ev_halo, cudaEventDisableTiming));
  }
  if (h.n_send) {
    k_pack<<<div_up(h.n_send, 256), 256, 0, stream()>>>(h.n_send, h.send_idx.p, x, xs, h.send_buf.p);
    count_launch();
  }
  B200_CUDA(cudaEventRecord(ev_packed, stream()));
  B200_CUDA(cudaStreamWaitEvent(c.comm_stream, ev_packed, 0));
  std::vector<const void *> sp;
  std::vector<void *> rp;
  std::vector<size_t> sb, rb;
  for (size_t k = 0; k < h.send_rank.size(); ++k) {
    sp.push_back(h.send_buf.p + h.send_off[k]);
    sb.push_back((size_t)(h.send_off[k + 1] - h.send_off[k]) * sizeof(double));
  }
  for (size_t k = 0; k < h.recv_rank.size(); ++k) {
    rp.push_back(A.x_halo.p + h.recv_off[k]);
    rb.push_back((size_t)(h.recv_off[k + 1] - h.recv_off[k]) * sizeof(double));
  }
  exchange_dev((int)h.send_rank.size(), h.send_rank.data(), sp.data(), sb.data(), (int)h.recv_rank.size(), h.recv_rank.data(),
               rp.data(), rb.data(), c.comm_stream);
  B200_CUDA(cudaEventRecord(ev_halo, c.comm_stream));
}
void halo_exchange_end(ParCsr &) { B200_CUDA(cudaStreamWaitEvent(stream(), ev_halo, 0)); }

void par_spmv(ParCsr &A, SpmvMode mode, const SpmvArgs &args) {
  const bool dist = ctx().nranks > 1;
  if (dist) halo_exchange_begin(A, args.x, args.xs);
  spmv(A.diag, &A.sell, A.plan, mode, args);
  if (!dist) return;
  halo_exchange_end(A);
  int nr = (int)A.offd_rows.n;
  if (A.offd.nnz == 0 || nr == 0) return;
  CsrView O = A.offd.view();
  int grid = div_up(nr, 128);
  switch (mode) {
    case SPMV_AX: k_offd_correct<SPMV_AX><<<grid, 128, 0, stream()>>>(nr, A.offd_rows.p, O, A.x_halo.p, args); break;
    case SPMV_RES: k_offd_correct<SPMV_RES><<<grid, 128, 0, stream()>>>(nr, A.offd_rows.p, O, A.x_halo.p, args); break;
    case SPMV_JAC: k_offd_correct<SPMV_JAC><<<grid, 128, 0, stream()>>>(nr, A.offd_rows.p, O, A.x_halo.p, args); break;
    case SPMV_ADD: k_offd_correct<SPMV_ADD><<<grid, 128, 0, stream()>>>(nr, A.offd_rows.p, O, A.x_halo.p, args); break;
    case SPMV_CHEB: k_offd_correct<SPMV_CHEB><<<grid, 128, 0, stream()>>>(nr, A.offd_rows.p, O, A.x_halo.p, args); break;
  }
  count_launch();
  B200_KERNEL_CHECK();
}

}  // namespace b200
