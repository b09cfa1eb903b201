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

// peer-memory arena (comm_peer.cu)
bool peer_active();
char *peer_base(int rank);
i64 peer_alloc(size_t bytes);
void peer_free(i64 off);

void HaloPlan::release() {
  if (region >= 0) peer_free(region);
  region = -1;
  p2p = false;
}
HaloPlan &HaloPlan::operator=(HaloPlan &&o) noexcept {
  if (this == &o) return *this;
  release();
  recv_rank = std::move(o.recv_rank); recv_off = std::move(o.recv_off);
  send_rank = std::move(o.send_rank); send_off = std::move(o.send_off);
  send_idx = std::move(o.send_idx); send_buf = std::move(o.send_buf);
  n_send = o.n_send;
  p2p = o.p2p; region = o.region; state = std::move(o.state); flags = o.flags; buf[0] = o.buf[0]; buf[1] = o.buf[1];
  send_off_dev = std::move(o.send_off_dev); push_dst = std::move(o.push_dst); ticket = std::move(o.ticket);
  o.region = -1; o.p2p = false; o.n_send = 0;
  return *this;
}

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
// ---- peer-memory exchange --------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// Pack x[idx[k]] (scaled by xs) and store it straight into the halo buffer of the rank that needs it (peer memory over
// NVLink); dst[3 s + parity] is the address of this rank's segment in destination s's buffer, dst[3 s + 2] its flag there.
// The last block to finish publishes `seq` in every destination's flag.
__global__ void k_pack_push(int n, const int *__restrict__ idx, const double *__restrict__ x, const double *__restrict__ xs, int nseg,
                            const int *__restrict__ soff, const unsigned long long *__restrict__ dst,
                            const unsigned long long *state, unsigned int *ticket) {
  __shared__ bool last;
  const unsigned long long seq = state[0] + 1;   // advanced by k_halo_wait, which follows this kernel in stream order
  const int parity = (int)(seq & 1);
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) {
    int s = 0;
    while (s + 1 < nseg && soff[s + 1] <= k) ++s;
    int i = idx[k];
    double v = x[i];
    if (xs) v *= xs[i];
    reinterpret_cast<double *>(dst[3 * s + parity])[k - soff[s]] = v;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (last) {
    __threadfence_system();
    for (int s = threadIdx.x; s < nseg; s += blockDim.x) st_release_sys(reinterpret_cast<unsigned long long *>(dst[3 * s + 2]), seq);
    if (threadIdx.x == 0) *ticket = 0;
  }
}
// Wait until every sending neighbour has published `seq` (its part of the halo is then complete in buffer seq & 1).
// A peer that never arrives trips the watchdog instead of hanging the device.
__global__ void k_halo_wait(int nflags, const unsigned long long *flags, unsigned long long *state, const double *buf0,
                            const double *buf1) {
  const int t = threadIdx.x;
  const unsigned long long seq = state[0] + 1;
  for (int k = t; k < nflags; k += blockDim.x) {
    long long t0 = clock64();
    while (ld_acquire_sys(flags + k) < seq) {
      if (clock64() - t0 > 40000000000LL) {  // ~20 s at 2 GHz
        printf("[b200hypre] halo exchange: neighbour %d did not deliver sequence %llu\n", k, seq);
        __trap();
      }
    }
  }
  __syncthreads();
  if (t == 0) {
    state[1] = (unsigned long long)((seq & 1) ? buf1 : buf0);
    state[0] = seq;
  }
}

__global__ void k_global_to_local(int n, const i64 *__restrict__ g, i64 start, int *__restrict__ out) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) out[k] = (int)(g[k] - start);
}

template <int MODE>
__global__ void k_offd_correct(int nrows, const int *__restrict__ rows, CsrView O, const double *const *xhp, SpmvArgs a) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nrows) return;
  const double *xh = *xhp;
  int i = rows[t];
  if (MODE == SPMV_JAC && a.cf != nullptr && a.cf[i] != a.pt) return;
  double s = 0.0;
  for (int k = O.ptr[i]; k < O.ptr[i + 1]; ++k) s += O.val[k] * __ldcg(xh + O.col[k]);  // halo may be peer-written: L2
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
  HaloPlan &h = A.halo;
  h.release();
  i64 mine[2] = {A.row_start, A.col_start};
  std::vector<i64> all(2 * np);
  allgather_host_i64(mine, all.data(), 2);   // also the barrier that makes arena regions released so far reusable
  for (int r = 0; r < np; ++r) { A.row_starts[r] = all[2 * r]; A.col_starts[r] = all[2 * r + 1]; }
  A.row_starts[np] = A.g_rows + A.row_starts[0];
  A.col_starts[np] = A.g_cols + A.col_starts[0];
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
  {
    unsigned long long st[2] = {0ULL, (unsigned long long)A.x_halo.p};
    h.state.from_host(st, 2);
  }
  if (peer_active()) {
    // region layout: flags (one per sending neighbour, padded to 256 B) | buffer 0 | buffer 1 (each padded to 256 B)
    auto layout = [](i64 nrecv, i64 nhalo, size_t &flag_bytes, size_t &buf_bytes) {
      flag_bytes = ((size_t)std::max<i64>(nrecv, 1) * 8 + 255) / 256 * 256;
      buf_bytes = ((size_t)std::max<i64>(nhalo, 1) * 8 + 255) / 256 * 256;
    };
    size_t fb, bb;
    layout((i64)h.recv_rank.size(), A.n_halo, fb, bb);
    i64 off = peer_alloc(fb + 2 * bb);
    std::vector<i64> offs(np);
    allgather_host_i64(&off, offs.data(), 1);
    bool ok = true;
    for (int r = 0; r < np; ++r) ok = ok && offs[r] >= 0;
    if (!ok) {
      peer_free(off);   // some rank ran out of arena: everybody keeps the NCCL path for this matrix
    } else {
      h.region = off;
      h.p2p = true;
      char *me = peer_base(c.rank) + off;
      h.flags = reinterpret_cast<unsigned long long *>(me);
      h.buf[0] = reinterpret_cast<double *>(me + fb);
      h.buf[1] = reinterpret_cast<double *>(me + fb + bb);
      B200_CUDA(cudaMemsetAsync(me, 0, fb + 2 * bb, stream()));
      // where my segment lives on every destination: its halo is ordered by owner rank, allreq[d][*] are its segment sizes
      std::vector<unsigned long long> dst(3 * std::max<size_t>(h.send_rank.size(), 1), 0);
      for (size_t k = 0; k < h.send_rank.size(); ++k) {
        int d = h.send_rank[k];
        i64 seg_off = 0, slot = 0, nrecv_d = 0, nhalo_d = 0;
        for (int r = 0; r < np; ++r) {
          i64 cnt = allreq[(size_t)d * np + r];
          if (cnt > 0) {
            if (r < c.rank) { seg_off += cnt; ++slot; }
            ++nrecv_d;
            nhalo_d += cnt;
          }
        }
        size_t fbd, bbd;
        layout(nrecv_d, nhalo_d, fbd, bbd);
        char *pd = peer_base(d) + offs[d];
        dst[3 * k + 0] = (unsigned long long)(pd + fbd + (size_t)seg_off * 8);
        dst[3 * k + 1] = (unsigned long long)(pd + fbd + bbd + (size_t)seg_off * 8);
        dst[3 * k + 2] = (unsigned long long)(pd + (size_t)slot * 8);
      }
      h.push_dst.from_host(dst);
      h.send_off_dev.from_host(h.send_off);
      h.ticket.alloc(1);
      h.ticket.zero();
    }
  }
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
  if (h.p2p) {  // push into the neighbours' halo buffers over NVLink; they wait on their flags (halo_exchange_end)
    if (!h.send_rank.empty()) {
      k_pack_push<<<div_up(std::max(h.n_send, 1), 256), 256, 0, stream()>>>(h.n_send, h.send_idx.p, x, xs, (int)h.send_rank.size(),
                                                                            h.send_off_dev.p, h.push_dst.p, h.state.p, h.ticket.p);
      count_launch();
    }
    return;
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
void halo_exchange_end(ParCsr &A) {
  HaloPlan &h = A.halo;
  if (h.p2p) {
    // always launched: it also advances the plan's sequence number and selects the buffer the consumers read
    k_halo_wait<<<1, 64, 0, stream()>>>((int)h.recv_rank.size(), h.flags, h.state.p, h.buf[0], h.buf[1]);
    count_launch();
    return;
  }
  B200_CUDA(cudaStreamWaitEvent(stream(), ev_halo, 0));
}

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
    case SPMV_AX: k_offd_correct<SPMV_AX><<<grid, 128, 0, stream()>>>(nr, A.offd_rows.p, O, A.xhp(), args); break;
    case SPMV_RES: k_offd_correct<SPMV_RES><<<grid, 128, 0, stream()>>>(nr, A.offd_rows.p, O, A.xhp(), args); break;
    case SPMV_JAC: k_offd_correct<SPMV_JAC><<<grid, 128, 0, stream()>>>(nr, A.offd_rows.p, O, A.xhp(), args); break;
    case SPMV_ADD: k_offd_correct<SPMV_ADD><<<grid, 128, 0, stream()>>>(nr, A.offd_rows.p, O, A.xhp(), args); break;
    case SPMV_CHEB: k_offd_correct<SPMV_CHEB><<<grid, 128, 0, stream()>>>(nr, A.offd_rows.p, O, A.xhp(), args); break;
  }
  count_launch();
  B200_KERNEL_CHECK();
}

}  // namespace b200
