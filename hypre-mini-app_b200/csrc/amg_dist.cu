// amg_dist.cu -- BoomerAMG setup on a row-partitioned ParCSR matrix (one rank per GPU): distributed strength, PMIS,
// coarse numbering, extended+i interpolation, restriction and the Galerkin product.
// Replaces the multi-rank path of HYPRE_BoomerAMGSetup (/root/reference/src/HypreSystem.cpp:323 via :692) on the block-row
// partition of /root/reference/src/HypreSystem.cpp:525-544; algorithm definitions SURVEY.md A.3-A.6, exchanges §2.5 / §8e.
//
// Approach: every phase works on an EXTENDED LOCAL matrix.  Own rows are kept in merged form (diag block entries, then
// offd block entries: hypre's ParCSR storage order); the external rows a phase needs (A/S rows of halo points for
// ext+i, P rows for A*P, Q rows for R*Q, the transposed P_offd for R) are fetched from their owners in the owner's
// storage order with GLOBAL column ids, columns are renumbered into  own | external (ascending global id), and then the
// single-GPU kernels of amg_setup.cu / amg_interp.cu / spgemm.cu run unchanged.  Because those kernels accumulate in
// row-storage order, the floating-point order -- and with it every discrete decision of the next level -- equals the
// oracle's N-virtual-rank run (oracle/oracle.cpp), for any number of ranks.
//
// Compiled with -fmad=false like the rest of the setup.
#include "amg.h"
#include "amg_rows.cuh"
#include <cub/cub.cuh>
#include <algorithm>

namespace b200 {

void build_partition_and_halo(ParCsr &A);
void allgather_host_i64(const i64 *in, i64 *out, int count);
void exchange_dev(int nsend, const int *send_rank, const void *const *send_ptr, const size_t *send_bytes, int nrecv,
                  const int *recv_rank, void *const *recv_ptr, const size_t *recv_bytes, cudaStream_t s);

namespace {

inline rows::Csr rv(const DCsr &A) { return rows::Csr{A.nr, A.nc, A.ptr.p, A.col.p, A.val.p}; }
inline int grid_for(i64 n) { return div_up(n, 256); }
#define DLAUNCH(kern, n, ...)                                                        \
  do {                                                                               \
    if ((n) > 0) { kern<<<grid_for(n), 256, 0, stream()>>>(__VA_ARGS__); count_launch(); } \
  } while (0)

template <class T>
T fetch1(const T *dev) {
  T h;
  B200_CUDA(cudaMemcpyAsync(&h, dev, sizeof(T), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  return h;
}
template <class T>
void exscan(const T *in, T *out, i64 n) {
  if (n <= 0) return;
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, in, out, (int)n, stream());
  DVec<char> tmp(tb);
  cub::DeviceScan::ExclusiveSum(tmp.p, tb, in, out, (int)n, stream());
  count_launch(2);
}
template <class T>
void dcopy(T *dst, const T *src, i64 n) {
  if (n > 0) B200_CUDA(cudaMemcpyAsync(dst, src, (size_t)n * sizeof(T), cudaMemcpyDeviceToDevice, stream()));
}

// sorted unique copy of keys[0..n): returns the count
int sort_unique(const i64 *keys, i64 n, DVec<i64> &out) {
  out.alloc((size_t)std::max<i64>(n, 1));
  if (n == 0) return 0;
  DVec<i64> sorted((size_t)n);
  DVec<int> nsel(1);
  size_t tb = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, tb, keys, sorted.p, (int)n, 0, 64, stream());
  DVec<char> tmp(tb);
  cub::DeviceRadixSort::SortKeys(tmp.p, tb, keys, sorted.p, (int)n, 0, 64, stream());
  size_t tb2 = 0;
  cub::DeviceSelect::Unique(nullptr, tb2, sorted.p, out.p, nsel.p, (int)n, stream());
  DVec<char> tmp2(tb2);
  cub::DeviceSelect::Unique(tmp2.p, tb2, sorted.p, out.p, nsel.p, (int)n, stream());
  count_launch(4);
  return fetch1(nsel.p);
}

__device__ __forceinline__ int lower_bound_dev(const i64 *a, int n, i64 key) {
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = lo + ((hi - lo) >> 1);
    if (a[mid] < key) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// ---- numbering:  own ids [0, n_own) <-> global [start, start + n_own);  external ids n_own + k <-> ext[k] (sorted) ----------
struct Numbering {
  i64 start;
  int n_own;
  const i64 *ext;   // first list (sorted)
  int n_ext;
  const i64 *ext2;  // optional second list (sorted), ids continue after the first
  int n_ext2;
};
__device__ __forceinline__ int to_local(const Numbering &N, i64 g) {
  if (g >= N.start && g < N.start + N.n_own) return (int)(g - N.start);
  int k = lower_bound_dev(N.ext, N.n_ext, g);
  if (k < N.n_ext && N.ext[k] == g) return N.n_own + k;
  if (N.n_ext2 > 0) {
    int k2 = lower_bound_dev(N.ext2, N.n_ext2, g);
    if (k2 < N.n_ext2 && N.ext2[k2] == g) return N.n_own + N.n_ext + k2;
  }
  return -1;
}
__device__ __forceinline__ i64 to_global(const Numbering &N, int c) {
  if (c < N.n_own) return N.start + c;
  c -= N.n_own;
  if (c < N.n_ext) return N.ext[c];
  return N.ext2[c - N.n_ext];
}

__global__ void k_map_to_local(i64 n, const i64 *g, Numbering N, int *out, int *bad) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  int c = to_local(N, g[k]);
  if (c < 0) atomicAdd(bad, 1);
  out[k] = c;
}
// flag = 1 for global ids that are neither own nor in the first external list
__global__ void k_flag_unknown(i64 n, const i64 *g, Numbering N, int *flag) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  flag[k] = (to_local(N, g[k]) < 0) ? 1 : 0;
  if (k == 0) flag[n] = 0;
}
__global__ void k_compact_i64(i64 n, const i64 *in, const int *flag, const int *pos, i64 *out) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n && flag[k]) out[pos[k]] = in[k];
}
// old external index (through old list) -> new local id
__global__ void k_remap_cols(i64 n, int *col, int n_own, const i64 *old_ext, Numbering N, int *bad) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  int c = col[k];
  if (c < n_own) return;
  int nc = to_local(N, old_ext[c - n_own]);
  if (nc < 0) atomicAdd(bad, 1);
  col[k] = nc;
}

// ids of `g` (count n) that are unknown to numbering N -> sorted unique list
int collect_unknown(const i64 *g, i64 n, const Numbering &N, DVec<i64> &out) {
  if (n == 0) { out.alloc(1); return 0; }
  DVec<int> flag((size_t)n + 1), pos((size_t)n + 1);
  DLAUNCH(k_flag_unknown, n, n, g, N, flag.p);
  exscan(flag.p, pos.p, n + 1);
  int cnt = fetch1(pos.p + n);
  DVec<i64> sel((size_t)std::max(cnt, 1));
  DLAUNCH(k_compact_i64, n, n, g, flag.p, pos.p, sel.p);
  return sort_unique(sel.p, cnt, out);
}

// ---- merged form of a ParCsr: row = diag entries then offd entries, offd column j -> n_cols_local + j ------------------------
__global__ void k_merge_count(int n, const int *dptr, const int *optr, int *cnt) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) cnt[i] = (dptr[i + 1] - dptr[i]) + (optr ? (optr[i + 1] - optr[i]) : 0);
  if (i == n) cnt[i] = 0;
}
// G lanes per row: coalesced copies (thread-per-row loops run at a fraction of the HBM rate on 30-60 entry rows)
template <int G>
__global__ void k_merge_fill(int n, CsrView D, CsrView O, int shift, const int *mptr, int *mcol, double *mval) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) / G, lane = threadIdx.x % G;
  if (i >= n) return;
  const int w = mptr[i], db = D.ptr[i], dn = D.ptr[i + 1] - db;
  for (int k = lane; k < dn; k += G) { mcol[w + k] = D.col[db + k]; mval[w + k] = D.val[db + k]; }
  if (O.ptr) {
    const int ob = O.ptr[i], on = O.ptr[i + 1] - ob;
    for (int k = lane; k < on; k += G) { mcol[w + dn + k] = O.col[ob + k] + shift; mval[w + dn + k] = O.val[ob + k]; }
  }
}
inline int lanes_for(i64 nnz, int n) {
  double mean = n > 0 ? (double)nnz / n : 1.0;
  return mean <= 10.0 ? 8 : 32;
}
void merge_parcsr(const ParCsr &A, DCsr &M) {
  int n = A.n;
  M.nr = n;
  M.nc = A.n_cols_local + A.n_halo;
  M.ptr.alloc((size_t)n + 1);
  DVec<int> cnt((size_t)n + 1);
  const bool has_offd = A.offd.nnz > 0;
  DLAUNCH(k_merge_count, (i64)n + 1, n, A.diag.ptr.p, has_offd ? A.offd.ptr.p : nullptr, cnt.p);
  exscan(cnt.p, M.ptr.p, (i64)n + 1);
  M.nnz = A.diag.nnz + (has_offd ? A.offd.nnz : 0);
  M.col.alloc((size_t)std::max<i64>(M.nnz, 1));
  M.val.alloc((size_t)std::max<i64>(M.nnz, 1));
  CsrView O = has_offd ? A.offd.view() : CsrView{0, 0, nullptr, nullptr, nullptr};
  if (lanes_for(M.nnz, n) == 8) DLAUNCH(k_merge_fill<8>, (i64)n * 8, n, A.diag.view(), O, A.n_cols_local, M.ptr.p, M.col.p, M.val.p);
  else DLAUNCH(k_merge_fill<32>, (i64)n * 32, n, A.diag.view(), O, A.n_cols_local, M.ptr.p, M.col.p, M.val.p);
  B200_KERNEL_CHECK();
}

// ---- split a merged matrix (columns own | external list) into the diag / offd blocks of a ParCsr; the offd column map is
// compressed to the external columns that are actually referenced ------------------------------------------------------------
template <int G>
__device__ __forceinline__ unsigned grp_mask() {
  if (G >= 32) return 0xffffffffu;
  return ((1u << (G & 31)) - 1u) << (((threadIdx.x & 31) / G) * G);
}
template <int G>
__global__ void k_split_count(int n, const int *mptr, const int *mcol, int n_own, int *dcnt, int *ocnt, int *used) {
  // collectives name the full warp (groups beyond the last row run predicated off): a per-group mask is serialised
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) / G, lane = threadIdx.x % G;
  int d = 0, o = 0;
  if (i < n)
    for (int k = mptr[i] + lane; k < mptr[i + 1]; k += G) {
      int c = mcol[k];
      if (c < n_own) ++d;
      else { ++o; used[c - n_own] = 1; }
    }
#pragma unroll
  for (int off = G >> 1; off > 0; off >>= 1) { d += __shfl_xor_sync(0xffffffffu, d, off, G); o += __shfl_xor_sync(0xffffffffu, o, off, G); }
  if (lane == 0 && i <= n) { dcnt[i] = d; ocnt[i] = o; }   // i == n: the terminating zero of the count arrays
}
// order-preserving split: positions inside the two output rows come from ballot prefixes over each chunk of G entries
template <int G>
__global__ void k_split_fill2(int n, const int *mptr, const int *mcol, const double *mval, int n_own, const int *newidx,
                              const int *dptr, int *dcol, double *dval, const int *optr, int *ocol, double *oval) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) / G, lane = threadIdx.x % G;
  const bool act = i < n;
  const unsigned gbits = (G >= 32) ? 0xffffffffu : ((1u << (G & 31)) - 1u);
  const int first = (threadIdx.x & 31) - lane;
  const unsigned lt = (1u << lane) - 1u;
  int wd = act ? dptr[i] : 0, wo = act ? optr[i] : 0;
  const int b = act ? mptr[i] : 0, e = act ? mptr[i + 1] : 0;
  const int longest = __reduce_max_sync(0xffffffffu, e - b);   // chunk loop uniform over the warp, full-warp ballots
  for (int off = 0; off < longest; off += G) {
    const int k = b + off + lane;
    int c = 0;
    bool valid = k < e, isd = false;
    if (valid) { c = mcol[k]; isd = c < n_own; }
    const unsigned bd = (__ballot_sync(0xffffffffu, valid && isd) >> first) & gbits;
    const unsigned bo = (__ballot_sync(0xffffffffu, valid && !isd) >> first) & gbits;
    if (valid) {
      if (isd) { int w = wd + __popc(bd & lt); dcol[w] = c; dval[w] = mval[k]; }
      else { int w = wo + __popc(bo & lt); ocol[w] = newidx[c - n_own]; oval[w] = mval[k]; }
    }
    wd += __popc(bd);
    wo += __popc(bo);
  }
}
__global__ void k_gather_used(int n, const int *used, const int *pos, const i64 *ext, i64 *out) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n && used[k]) out[pos[k]] = ext[k];
}
__global__ void k_flag_nonempty(int n, const int *ptr, int *flag) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flag[i] = (ptr[i + 1] > ptr[i]) ? 1 : 0;
  if (i == n) flag[i] = 0;
}
__global__ void k_compact_idx(int n, const int *flag, const int *pos, int *list) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && flag[i]) list[pos[i]] = i;
}
void build_offd_rows(ParCsr &A) {
  int n = A.n;
  A.offd_rows.release();
  if (A.offd.nnz == 0 || n == 0) return;
  DVec<int> flag((size_t)n + 1), pos((size_t)n + 1);
  DLAUNCH(k_flag_nonempty, (i64)n + 1, n, A.offd.ptr.p, flag.p);
  exscan(flag.p, pos.p, (i64)n + 1);
  int nr = fetch1(pos.p + n);
  A.offd_rows.alloc((size_t)nr);
  DLAUNCH(k_compact_idx, n, n, flag.p, pos.p, A.offd_rows.p);
}
// M: n rows, columns [0, n_own) own, n_own + k -> ext[k]; fills out.diag / out.offd / out.col_map_offd / out.n_halo / offd_rows
void split_merged(const DCsr &M, int n_own, const i64 *ext, int n_ext, ParCsr &out) {
  int n = M.nr;
  out.n = n;
  out.n_cols_local = n_own;
  DVec<int> dcnt((size_t)n + 1), ocnt((size_t)n + 1), used((size_t)n_ext + 1), pos((size_t)n_ext + 1);
  used.zero();
  const int G = lanes_for(M.nnz, n);
  if (G == 8) DLAUNCH(k_split_count<8>, ((i64)n + 1) * 8, n, M.ptr.p, M.col.p, n_own, dcnt.p, ocnt.p, used.p);
  else DLAUNCH(k_split_count<32>, ((i64)n + 1) * 32, n, M.ptr.p, M.col.p, n_own, dcnt.p, ocnt.p, used.p);
  out.diag.nr = out.offd.nr = n;
  out.diag.nc = n_own;
  out.diag.ptr.alloc((size_t)n + 1);
  out.offd.ptr.alloc((size_t)n + 1);
  exscan(dcnt.p, out.diag.ptr.p, (i64)n + 1);
  exscan(ocnt.p, out.offd.ptr.p, (i64)n + 1);
  exscan(used.p, pos.p, (i64)n_ext + 1);
  out.diag.nnz = fetch1(out.diag.ptr.p + n);
  out.offd.nnz = fetch1(out.offd.ptr.p + n);
  int nh = fetch1(pos.p + n_ext);
  out.n_halo = nh;
  out.offd.nc = nh;
  DVec<i64> cmap((size_t)std::max(nh, 1));
  DLAUNCH(k_gather_used, n_ext, n_ext, used.p, pos.p, ext, cmap.p);
  out.col_map_offd.resize(nh);
  if (nh) {
    B200_CUDA(cudaMemcpyAsync(out.col_map_offd.data(), cmap.p, (size_t)nh * sizeof(i64), cudaMemcpyDeviceToHost, stream()));
    B200_CUDA(cudaStreamSynchronize(stream()));
  }
  out.diag.col.alloc((size_t)std::max<i64>(out.diag.nnz, 1));
  out.diag.val.alloc((size_t)std::max<i64>(out.diag.nnz, 1));
  out.offd.col.alloc((size_t)std::max<i64>(out.offd.nnz, 1));
  out.offd.val.alloc((size_t)std::max<i64>(out.offd.nnz, 1));
  if (G == 8)
    DLAUNCH(k_split_fill2<8>, (i64)n * 8, n, M.ptr.p, M.col.p, M.val.p, n_own, pos.p, out.diag.ptr.p, out.diag.col.p, out.diag.val.p,
            out.offd.ptr.p, out.offd.col.p, out.offd.val.p);
  else
    DLAUNCH(k_split_fill2<32>, (i64)n * 32, n, M.ptr.p, M.col.p, M.val.p, n_own, pos.p, out.diag.ptr.p, out.diag.col.p, out.diag.val.p,
            out.offd.ptr.p, out.offd.col.p, out.offd.val.p);
  build_offd_rows(out);
  B200_KERNEL_CHECK();
}

// ---- fixed-size halo exchanges on the plan of a ParCsr (4- or 8-byte elements) -----------------------------------------------
template <class T>
__global__ void k_pack_t(int n, const int *__restrict__ idx, const T *__restrict__ x, T *__restrict__ out) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) out[k] = x[idx[k]];
}
// halo[j] = value of the owner's entry col_map_offd[j]
template <class T>
void halo_fwd(ParCsr &A, const T *own, T *halo) {
  HaloPlan &h = A.halo;
  DVec<T> sbuf((size_t)std::max(h.n_send, 1));
  if (h.n_send) { k_pack_t<T><<<grid_for(h.n_send), 256, 0, stream()>>>(h.n_send, h.send_idx.p, own, sbuf.p); count_launch(); }
  std::vector<const void *> sp;
  std::vector<void *> rp;
  std::vector<size_t> sb, rb;
  for (size_t k = 0; k < h.send_rank.size(); ++k) {
    sp.push_back(sbuf.p + h.send_off[k]);
    sb.push_back((size_t)(h.send_off[k + 1] - h.send_off[k]) * sizeof(T));
  }
  for (size_t k = 0; k < h.recv_rank.size(); ++k) {
    rp.push_back(halo + h.recv_off[k]);
    rb.push_back((size_t)(h.recv_off[k + 1] - h.recv_off[k]) * sizeof(T));
  }
  exchange_dev((int)h.send_rank.size(), h.send_rank.data(), sp.data(), sb.data(), (int)h.recv_rank.size(), h.recv_rank.data(),
               rp.data(), rb.data(), stream());
  // no host synchronisation: the exchange is enqueued on stream() and sbuf is returned to the stream-ordered pool on the same
  // stream, so it cannot be reused before the send has read it
}
// reverse direction: every rank sends its halo values back to the owners; back[t] is the value a requester holds for the
// owner's entry send_idx[t] (t in send-list order, one slot per (requester, entry))
template <class T>
void halo_rev(ParCsr &A, const T *halo, T *back) {
  HaloPlan &h = A.halo;
  std::vector<const void *> sp;
  std::vector<void *> rp;
  std::vector<size_t> sb, rb;
  for (size_t k = 0; k < h.recv_rank.size(); ++k) {
    sp.push_back(halo + h.recv_off[k]);
    sb.push_back((size_t)(h.recv_off[k + 1] - h.recv_off[k]) * sizeof(T));
  }
  for (size_t k = 0; k < h.send_rank.size(); ++k) {
    rp.push_back(back + h.send_off[k]);
    rb.push_back((size_t)(h.send_off[k + 1] - h.send_off[k]) * sizeof(T));
  }
  exchange_dev((int)h.recv_rank.size(), h.recv_rank.data(), sp.data(), sb.data(), (int)h.send_rank.size(), h.send_rank.data(),
               rp.data(), rb.data(), stream());
}

// ---- variable-length row exchange ---------------------------------------------------------------------------------------------
struct RowsMsg {  // CSR of received rows: global column ids + one 8-byte payload per entry (double bits or an integer code)
  int nr = 0;
  i64 nnz = 0;
  DVec<int> ptr;
  DVec<i64> col, val;
};
__global__ void k_row_lens(int n, const int *ptr, int *len) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) len[i] = ptr[i + 1] - ptr[i];
}
// Sends rows [soff[k], soff[k+1]) of the CSR (sptr, scol, sval) to srank[k]; receives rows [roff[k], roff[k+1]) from rrank[k].
void exchange_rows(const std::vector<int> &srank, const std::vector<int> &soff, const int *sptr, const i64 *scol, const i64 *sval,
                   const std::vector<int> &rrank, const std::vector<int> &roff, RowsMsg &out) {
  const int ns = soff.empty() ? 0 : soff.back(), nr = roff.empty() ? 0 : roff.back();
  DVec<int> slen((size_t)std::max(ns, 1)), rlen((size_t)nr + 1);
  rlen.zero();
  DLAUNCH(k_row_lens, ns, ns, sptr, slen.p);
  std::vector<const void *> sp;
  std::vector<void *> rp;
  std::vector<size_t> sb, rb;
  for (size_t k = 0; k < srank.size(); ++k) { sp.push_back(slen.p + soff[k]); sb.push_back((size_t)(soff[k + 1] - soff[k]) * sizeof(int)); }
  for (size_t k = 0; k < rrank.size(); ++k) { rp.push_back(rlen.p + roff[k]); rb.push_back((size_t)(roff[k + 1] - roff[k]) * sizeof(int)); }
  exchange_dev((int)srank.size(), srank.data(), sp.data(), sb.data(), (int)rrank.size(), rrank.data(), rp.data(), rb.data(), stream());
  out.nr = nr;
  out.ptr.alloc((size_t)nr + 1);
  exscan(rlen.p, out.ptr.p, (i64)nr + 1);
  std::vector<int> hs((size_t)ns + 1, 0), hr((size_t)nr + 1, 0);
  if (ns) B200_CUDA(cudaMemcpyAsync(hs.data(), sptr, ((size_t)ns + 1) * sizeof(int), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaMemcpyAsync(hr.data(), out.ptr.p, ((size_t)nr + 1) * sizeof(int), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  out.nnz = hr[nr];
  out.col.alloc((size_t)std::max<i64>(out.nnz, 1));
  out.val.alloc((size_t)std::max<i64>(out.nnz, 1));
  for (int pass = 0; pass < 2; ++pass) {
    const i64 *src = pass == 0 ? scol : sval;
    i64 *dst = pass == 0 ? out.col.p : out.val.p;
    sp.clear(); rp.clear(); sb.clear(); rb.clear();
    for (size_t k = 0; k < srank.size(); ++k) {
      sp.push_back(src + (hs[soff[k]] - hs[0]));
      sb.push_back((size_t)(hs[soff[k + 1]] - hs[soff[k]]) * sizeof(i64));
    }
    for (size_t k = 0; k < rrank.size(); ++k) {
      rp.push_back(dst + hr[roff[k]]);
      rb.push_back((size_t)(hr[roff[k + 1]] - hr[roff[k]]) * sizeof(i64));
    }
    exchange_dev((int)srank.size(), srank.data(), sp.data(), sb.data(), (int)rrank.size(), rrank.data(), rp.data(), rb.data(), stream());
  }
  // the payload exchanges stay enqueued on stream(): consumers are stream-ordered behind them
}

// rows `list[0..nrows)` of a local matrix in (own | external) numbering -> packed CSR with global columns.  Payload: the
// matrix values (val != nullptr) or an integer code looked up per column (code_of[c], c = local column id).
__global__ void k_pack_rows_count(int nrows, const int *list, const int *mptr, int *cnt) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nrows) { int i = list[t]; cnt[t] = mptr[i + 1] - mptr[i]; }
  if (t == nrows) cnt[t] = 0;
}
__global__ void k_pack_rows_fill(int nrows, const int *list, const int *mptr, const int *mcol, const double *mval, const i64 *code_of,
                                 Numbering N, const int *sptr, i64 *scol, i64 *sval) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nrows) return;
  int i = list[t];
  int w = sptr[t];
  for (int k = mptr[i]; k < mptr[i + 1]; ++k, ++w) {
    int c = mcol[k];
    scol[w] = to_global(N, c);
    sval[w] = mval ? __double_as_longlong(mval[k]) : code_of[c];
  }
}
struct Packed {
  DVec<int> ptr;
  DVec<i64> col, val;
};
void pack_rows(int nrows, const int *list, const int *mptr, const int *mcol, const double *mval, const i64 *code_of,
               const Numbering &N, Packed &P) {
  DVec<int> cnt((size_t)nrows + 1);
  P.ptr.alloc((size_t)nrows + 1);
  DLAUNCH(k_pack_rows_count, (i64)nrows + 1, nrows, list, mptr, cnt.p);
  exscan(cnt.p, P.ptr.p, (i64)nrows + 1);
  int nnz = fetch1(P.ptr.p + nrows);
  P.col.alloc((size_t)std::max(nnz, 1));
  P.val.alloc((size_t)std::max(nnz, 1));
  DLAUNCH(k_pack_rows_fill, nrows, nrows, list, mptr, mcol, mval, code_of, N, P.ptr.p, P.col.p, P.val.p);
  B200_KERNEL_CHECK();
}
// fetch the rows of the halo points of H (H's column halo) from their owners
void fetch_halo_rows(ParCsr &H, const int *mptr, const int *mcol, const double *mval, const i64 *code_of, const Numbering &N,
                     RowsMsg &out) {
  HaloPlan &h = H.halo;
  Packed P;
  pack_rows(h.n_send, h.send_idx.p, mptr, mcol, mval, code_of, N, P);
  exchange_rows(h.send_rank, h.send_off, P.ptr.p, P.col.p, P.val.p, h.recv_rank, h.recv_off, out);
}

// stack  [own rows (local columns) ; received rows (global columns -> local through N)]  into one CSR
__global__ void k_shift_ptr(int n, const int *in, int shift, int *out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[i] + shift;
}
__global__ void k_bits_to_double(i64 n, const i64 *in, double *out) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) out[k] = __longlong_as_double(in[k]);
}
void stack_rows(int n_own_rows, const int *optr, const int *ocol, const double *oval, i64 onnz, const RowsMsg &ext, const Numbering &N,
                int ncols, bool with_values, DCsr &out) {
  out.nr = n_own_rows + ext.nr;
  out.nc = ncols;
  out.nnz = onnz + ext.nnz;
  B200_REQUIRE(out.nnz < (1LL << 31) - 8, "extended matrix exceeds 2^31 entries");
  out.ptr.alloc((size_t)out.nr + 1);
  out.col.alloc((size_t)std::max<i64>(out.nnz, 1));
  if (with_values) out.val.alloc((size_t)std::max<i64>(out.nnz, 1));
  dcopy(out.ptr.p, optr, (i64)n_own_rows + 1);
  if (n_own_rows == 0) B200_CUDA(cudaMemsetAsync(out.ptr.p, 0, sizeof(int), stream()));
  DLAUNCH(k_shift_ptr, ext.nr, ext.nr, ext.ptr.p + 1, (int)onnz, out.ptr.p + n_own_rows + 1);
  dcopy(out.col.p, ocol, onnz);
  if (with_values) dcopy(out.val.p, oval, onnz);
  DVec<int> bad(1);
  bad.zero();
  DLAUNCH(k_map_to_local, ext.nnz, ext.nnz, ext.col.p, N, out.col.p + onnz, bad.p);
  if (with_values) DLAUNCH(k_bits_to_double, ext.nnz, ext.nnz, ext.val.p, out.val.p + onnz);
  B200_REQUIRE(fetch1(bad.p) == 0, "extended matrix: a fetched column id is missing from the numbering");
  B200_KERNEL_CHECK();
}

// ---- distributed PMIS kernels (arrays indexed by merged ids: own [0, n), halo [n, n + nh)) ------------------------------------
__global__ void k_strength_count_d(rows::Csr A, double theta, double mrs, int *cnt) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < A.nr) cnt[i] = rows::strength_row(A, i, theta, mrs, nullptr);
  if (i == A.nr) cnt[i] = 0;
}
__global__ void k_strength_fill_d(rows::Csr A, double theta, double mrs, const int *Sptr, int *Scol) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < A.nr && Sptr[i + 1] > Sptr[i]) rows::strength_row(A, i, theta, mrs, Scol + Sptr[i]);
}
__global__ void k_col_count_d(i64 nnz, const int *Scol, int *cnt) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nnz) atomicAdd(cnt + Scol[k], 1);
}
__global__ void k_scatter_add_int(int n, const int *idx, const int *v, int *out) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n && v[t]) atomicAdd(out + idx[t], v[t]);
}
__global__ void k_scatter_min_int(int n, const int *idx, const int *v, int *out) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) atomicMin(out + idx[t], v[t]);
}
__global__ void k_pmis_init_d(int n, const int *Sptr, const int *cnt, int seed, double *measure, int *cf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (Sptr[i + 1] == Sptr[i]) { cf[i] = -3; measure[i] = 0.0; return; }
  double m = (double)cnt[i] + rows::hypre_rand_at(seed, i);
  if (m < 1.0) { cf[i] = -1; measure[i] = 0.0; }
  else { cf[i] = 0; measure[i] = m; }
}
// undecided points of this rank, added to a double (exact for counts below 2^53) so that the sum over ranks is one device
// all-reduce followed by a single read-back: one host synchronisation per sweep
__global__ void k_count_undecided_d(int n, const int *cf, double *counter) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int u = (i < n && cf[i] == 0) ? 1 : 0;
  unsigned b = __ballot_sync(0xffffffffu, u);
  if ((threadIdx.x & 31) == 0 && b) atomicAdd(counter, (double)__popc(b));
}
__global__ void k_is_init_d(int n, int nh, const int *cf, const double *measure, int *is) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) is[i] = (cf[i] == 0 && measure[i] > 1.0) ? 1 : 0;
  else if (i < n + nh) is[i] = 1;  // knock-out flag of a halo point, returned to its owner
}
__global__ void k_is_knock_d(int n, const int *Sptr, const int *Scol, const double *measure, const int *cf, int *is) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || cf[i] != 0) return;
  double mi = measure[i];
  if (!(mi > 1.0)) return;
  for (int k = Sptr[i]; k < Sptr[i + 1]; ++k) {
    int j = Scol[k];
    double mj = measure[j];
    if (mj > 1.0) {
      if (mi > mj) is[j] = 0;
      else if (mj > mi) is[i] = 0;
    }
  }
}
__global__ void k_set_c_d(int n, const int *is, int *cf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && cf[i] == 0 && is[i]) cf[i] = 1;
}
__global__ void k_set_f_d(int n, const int *Sptr, const int *Scol, int *cf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || cf[i] != 0) return;
  for (int k = Sptr[i]; k < Sptr[i + 1]; ++k)
    if (cf[Scol[k]] > 0) { cf[i] = -1; return; }
}
__global__ void k_clear_measure_d(int n, const int *cf, double *measure) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && cf[i] != 0) measure[i] = 0.0;
}
__global__ void k_c_flags_d(int n, const int *cf, int *flag) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flag[i] = cf[i] > 0 ? 1 : 0;
  if (i == n) flag[i] = 0;
}
__global__ void k_fold_sf_d(int n, const int *cf_raw, int *cf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) cf[i] = (cf_raw[i] == -3) ? -1 : cf_raw[i];
}
// code of a point: global coarse id when C, else its (negative) CF marker
__global__ void k_codes(int n, const int *cf, const int *f2c, i64 coarse_start, i64 *code) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) code[i] = (cf[i] > 0) ? coarse_start + f2c[i] : (i64)cf[i];
}
__global__ void k_fill_i64(i64 n, i64 *a, i64 v) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) a[k] = v;
}
// S rows of halo points carry (global column, code): record the code of every extended point they mention
__global__ void k_scatter_codes(i64 nnz, const int *col_local, const i64 *code_in, i64 *code) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nnz) code[col_local[k]] = code_in[k];
}
__global__ void k_cf_from_codes(int n, const i64 *code, int *cf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) cf[i] = code[i] >= 0 ? 1 : (int)code[i];
}
__global__ void k_ext_c_flags(int n_own, int n_tot, const i64 *code, int *flag) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_tot) flag[i] = (i >= n_own && code[i] >= 0) ? 1 : 0;
  if (i == n_tot) flag[i] = 0;
}
__global__ void k_f2c_ext(int n_own, int n_tot, const i64 *code, const int *f2c_own, int nc_own, const i64 *xc, int nxc, int *f2c) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_tot) return;
  if (i < n_own) { f2c[i] = f2c_own[i]; return; }
  i64 c = code[i];
  if (c < 0) { f2c[i] = -1; return; }
  f2c[i] = nc_own + lower_bound_dev(xc, nxc, c);
}
__global__ void k_expand_rows_d(int nr, const int *ptr, const int *rowmap, int *rowidx) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nr) return;
  int r = rowmap ? rowmap[i] : i;
  for (int k = ptr[i]; k < ptr[i + 1]; ++k) rowidx[k] = r;
}
__global__ void k_iota_d(i64 n, int *a) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) a[k] = (int)k;
}
__global__ void k_hist_d(i64 n, const int *key, int *cnt) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) atomicAdd(cnt + key[k], 1);
}
__global__ void k_gather_entries(i64 n, const int *perm, const int *col_in, const i64 *val_bits, int *col_out, double *val_out) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  int s = perm[k];
  col_out[k] = col_in[s];
  val_out[k] = __longlong_as_double(val_bits[s]);
}
__global__ void k_local_to_global_fine(i64 n, const int *col, i64 start, i64 *out) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) out[k] = start + col[k];
}
__global__ void k_double_to_bits(i64 n, const double *in, i64 *out) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) out[k] = __double_as_longlong(in[k]);
}
__global__ void k_smoother_diag_d(int n, CsrView D, CsrView O, int relax_type, double *d) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b = D.ptr[i], e = D.ptr[i + 1];
  if (b == e) { d[i] = 1.0; return; }
  double dg = D.val[b];
  if (relax_type == 7) { d[i] = dg; return; }
  double s = 0.0;
  for (int k = b; k < e; ++k) s = s + fabs(D.val[k]);
  if (O.ptr)
    for (int k = O.ptr[i]; k < O.ptr[i + 1]; ++k) s = s + fabs(O.val[k]);
  d[i] = (dg < 0) ? -s : s;
}

struct PhaseTimerD {
  Amg &amg;
  const char *name;
  cudaEvent_t e0, e1;
  PhaseTimerD(Amg &a, const char *n) : amg(a), name(n) {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, stream());
  }
  ~PhaseTimerD() {
    cudaEventRecord(e1, stream());
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    amg.setup_phase_ms[name] += ms;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
  }
};

void set_par_shape(ParCsr &M, int n, int ncl, i64 row_start, i64 col_start, i64 g_rows, i64 g_cols) {
  M.n = n;
  M.n_cols_local = ncl;
  M.row_start = row_start;
  M.col_start = col_start;
  M.g_rows = g_rows;
  M.g_cols = g_cols;
}

}  // namespace

// l1 norms / diagonal of a distributed level operator (row storage order: diag block, then offd block)
void smoother_diag_dist(ParCsr &A, int relax_type, double *d) {
  CsrView O = A.offd.nnz > 0 ? A.offd.view() : CsrView{0, 0, nullptr, nullptr, nullptr};
  DLAUNCH(k_smoother_diag_d, A.n, A.n, A.diag.view(), O, relax_type, d);
  B200_KERNEL_CHECK();
}

// One coarsening step of a distributed level.  Returns false when the level cannot be coarsened further.  Collective.
bool coarsen_level_dist(Amg &amg, AmgLevel &L, AmgLevel &Lc) {
  const AmgOptions &o = amg.opt;
  Context &c = ctx();
  ParCsr &Ap = *L.A;
  const int n = Ap.n, nh = Ap.n_halo, nm = n + nh;
  const int np = c.nranks;
  DVec<i64> a_colmap;  // device copy of A's halo column map
  a_colmap.from_host(Ap.col_map_offd.data(), Ap.col_map_offd.size());
  if (nh == 0) a_colmap.alloc(1);
  DCsr Am;
  {
    PhaseTimerD t0(amg, "merge");
    merge_parcsr(Ap, Am);
  }
  rows::Csr am = rv(Am);
  // ---- strength on the merged rows -----------------------------------------------------------------------------------------
  {
    PhaseTimerD t(amg, "strength");
    DVec<int> cnt((size_t)n + 1);
    DLAUNCH(k_strength_count_d, (i64)n + 1, am, o.strong_threshold, o.max_row_sum, cnt.p);
    L.Sptr.alloc((size_t)n + 1);
    exscan(cnt.p, L.Sptr.p, (i64)n + 1);
    L.S_nnz = fetch1(L.Sptr.p + n);
    L.Scol.alloc((size_t)std::max<i64>(L.S_nnz, 1));
    DLAUNCH(k_strength_fill_d, n, am, o.strong_threshold, o.max_row_sum, L.Sptr.p, L.Scol.p);
    B200_KERNEL_CHECK();
  }
  // ---- PMIS -------------------------------------------------------------------------------------------------------------------
  DVec<int> cfm((size_t)nm + 1);   // CF marker of own + halo points
  DVec<int> f2c((size_t)n + 1);
  int nc = 0;
  {
    PhaseTimerD t(amg, "pmis");
    DVec<int> cnt((size_t)nm + 1), back((size_t)std::max(Ap.halo.n_send, 1));
    cnt.zero();
    DLAUNCH(k_col_count_d, L.S_nnz, L.S_nnz, L.Scol.p, cnt.p);
    halo_rev<int>(Ap, cnt.p + n, back.p);   // dependencies counted on other ranks come home
    DLAUNCH(k_scatter_add_int, Ap.halo.n_send, Ap.halo.n_send, Ap.halo.send_idx.p, back.p, cnt.p);
    DVec<double> measure((size_t)nm + 1);
    DLAUNCH(k_pmis_init_d, n, n, L.Sptr.p, cnt.p, 2747 + c.rank, measure.p, cfm.p);
    DVec<int> is((size_t)nm + 1);
    DVec<double> counter(1);
    for (int sweep = 0; sweep < 100000; ++sweep) {
      counter.zero();
      DLAUNCH(k_count_undecided_d, n, n, cfm.p, counter.p);
      allreduce_sum_dev(counter.p, 1);
      if (fetch1(counter.p) == 0.0) break;
      halo_fwd<double>(Ap, measure.p, measure.p + n);
      DLAUNCH(k_is_init_d, nm, n, nh, cfm.p, measure.p, is.p);
      DLAUNCH(k_is_knock_d, n, n, L.Sptr.p, L.Scol.p, measure.p, cfm.p, is.p);
      halo_rev<int>(Ap, is.p + n, back.p);
      DLAUNCH(k_scatter_min_int, Ap.halo.n_send, Ap.halo.n_send, Ap.halo.send_idx.p, back.p, is.p);
      DLAUNCH(k_set_c_d, n, n, is.p, cfm.p);
      halo_fwd<int>(Ap, cfm.p, cfm.p + n);
      DLAUNCH(k_set_f_d, n, n, L.Sptr.p, L.Scol.p, cfm.p);
      DLAUNCH(k_clear_measure_d, n, n, cfm.p, measure.p);
    }
    halo_fwd<int>(Ap, cfm.p, cfm.p + n);
    B200_KERNEL_CHECK();
    DVec<int> flag((size_t)n + 1);
    DLAUNCH(k_c_flags_d, (i64)n + 1, n, cfm.p, flag.p);
    exscan(flag.p, f2c.p, (i64)n + 1);
    nc = fetch1(f2c.p + n);
    L.cf_raw.alloc((size_t)std::max(n, 1));
    dcopy(L.cf_raw.p, cfm.p, n);
  }
  // coarse partition
  std::vector<i64> ncs(np), cstarts(np + 1, 0);
  i64 mine = nc;
  allgather_host_i64(&mine, ncs.data(), 1);
  for (int r = 0; r < np; ++r) cstarts[r + 1] = cstarts[r] + ncs[r];
  const i64 gnc = cstarts[np], gn = Ap.g_rows, cstart = cstarts[c.rank];
  if (gnc == 0 || gnc == gn || gnc < o.min_coarse_size) return false;

  // ---- interpolation -------------------------------------------------------------------------------------------------------------
  DCsr Pm;            // merged P: columns own coarse | xc0 (sorted global coarse ids of the external C points seen)
  DVec<i64> xc0;
  int nxc0 = 0;
  {
    PhaseTimerD t(amg, "interp");
    // codes (global coarse id or negative marker) of own + halo points
    std::unique_ptr<PhaseTimerD> st(new PhaseTimerD(amg, "interp.fetch_rows"));
    DVec<i64> code((size_t)nm + 1);
    DLAUNCH(k_codes, n, n, cfm.p, f2c.p, cstart, code.p);
    halo_fwd<i64>(Ap, code.p, code.p + n);
    // A and S rows of the halo points, from their owners
    Numbering Nh{Ap.col_start, n, a_colmap.p, nh, nullptr, 0};
    RowsMsg extA, extS;
    fetch_halo_rows(Ap, Am.ptr.p, Am.col.p, Am.val.p, nullptr, Nh, extA);
    fetch_halo_rows(Ap, L.Sptr.p, L.Scol.p, nullptr, code.p, Nh, extS);
    // distance-2 points: columns of the fetched rows that are neither own nor halo
    st.reset();
    st.reset(new PhaseTimerD(amg, "interp.extend"));
    DVec<i64> newg;
    int nnew = collect_unknown(extA.col.p, extA.nnz, Nh, newg);
    const int ntot = nm + nnew;
    Numbering N2{Ap.col_start, n, a_colmap.p, nh, newg.p, nnew};
    DCsr Aext, Sext;
    stack_rows(n, Am.ptr.p, Am.col.p, Am.val.p, Am.nnz, extA, N2, ntot, true, Aext);
    stack_rows(n, L.Sptr.p, L.Scol.p, nullptr, L.S_nnz, extS, N2, ntot, false, Sext);
    // codes / CF markers / coarse columns of all extended points
    DVec<i64> code_ext((size_t)ntot + 1);
    DLAUNCH(k_fill_i64, (i64)ntot + 1, (i64)ntot + 1, code_ext.p, (i64)-1);
    dcopy(code_ext.p, code.p, nm);
    DLAUNCH(k_scatter_codes, extS.nnz, extS.nnz, Sext.col.p + L.S_nnz, extS.val.p, code_ext.p);
    DVec<int> cf_ext((size_t)ntot + 1), f2c_ext((size_t)ntot + 1);
    DLAUNCH(k_cf_from_codes, ntot, ntot, code_ext.p, cf_ext.p);
    dcopy(cf_ext.p, cfm.p, nm);  // own + halo markers keep -3 where it applies (codes carry it too)
    {
      DVec<int> flag((size_t)ntot + 1), pos((size_t)ntot + 1);
      DLAUNCH(k_ext_c_flags, (i64)ntot + 1, n, ntot, code_ext.p, flag.p);
      exscan(flag.p, pos.p, (i64)ntot + 1);
      int cnt = fetch1(pos.p + ntot);
      DVec<i64> sel((size_t)std::max(cnt, 1));
      DLAUNCH(k_compact_i64, ntot, ntot, code_ext.p, flag.p, pos.p, sel.p);
      nxc0 = sort_unique(sel.p, cnt, xc0);
    }
    DLAUNCH(k_f2c_ext, ntot, n, ntot, code_ext.p, f2c.p, nc, xc0.p, nxc0, f2c_ext.p);
    Pm.nc = nc + nxc0;
    st.reset();
    {
      PhaseTimerD t2(amg, "interp.build");
      if (o.interp_type == 0) build_interp_classical(Aext, Sext.ptr.p, Sext.col.p, Sext.nnz, cf_ext.p, f2c_ext.p, n, n, o.trunc_factor, o.p_max_elmts, Pm);
      else build_interp(Aext, Sext.ptr.p, Sext.col.p, Sext.nnz, cf_ext.p, f2c_ext.p, n, n, o.trunc_factor, o.p_max_elmts, Pm);
    }
    {
      PhaseTimerD t3(amg, "interp.split");
      split_merged(Pm, nc, xc0.p, nxc0, L.P);
    }
    set_par_shape(L.P, n, nc, Ap.row_start, cstart, gn, gnc);
    {
      PhaseTimerD t4(amg, "interp.halo_plan");
      build_partition_and_halo(L.P);
    }
    L.cf.alloc((size_t)std::max(n, 1));
    DLAUNCH(k_fold_sf_d, n, n, L.cf_raw.p, L.cf.p);
    B200_KERNEL_CHECK();
  }
  // ---- R = P^T, Q = A P, A_c = R Q ------------------------------------------------------------------------------------------------
  {
    PhaseTimerD t(amg, "rap");
    ParCsr &P = L.P, &R = L.R;
    // restriction: own part by local transposition, the rest arrives from the ranks whose P_offd references my coarse points
    std::unique_ptr<PhaseTimerD> sub(new PhaseTimerD(amg, "rap.transpose"));
    csr_transpose(P.diag, R.diag);
    {
      DCsr T;  // (P_offd)^T: one row per halo coarse column, entries = (local fine row ascending, value)
      csr_transpose(P.offd, T);
      DVec<i64> tcol((size_t)std::max<i64>(T.nnz, 1)), tval((size_t)std::max<i64>(T.nnz, 1));
      DLAUNCH(k_local_to_global_fine, T.nnz, T.nnz, T.col.p, Ap.row_start, tcol.p);
      DLAUNCH(k_double_to_bits, T.nnz, T.nnz, T.val.p, tval.p);
      RowsMsg in;  // row t: contributions to my coarse point send_idx[t] from one requester
      exchange_rows(P.halo.recv_rank, P.halo.recv_off, T.ptr.p, tcol.p, tval.p, P.halo.send_rank, P.halo.send_off, in);
      // COO (coarse row, global fine col) in arrival order = (rank ascending, fine ascending) -> stable sort by row
      DVec<i64> rmap;
      int nhr = sort_unique(in.col.p, in.nnz, rmap);
      Numbering Nr{Ap.row_start, n, rmap.p, nhr, nullptr, 0};
      DVec<int> lcol((size_t)std::max<i64>(in.nnz, 1)), rowkey((size_t)std::max<i64>(in.nnz, 1)), bad(1);
      bad.zero();
      DLAUNCH(k_map_to_local, in.nnz, in.nnz, in.col.p, Nr, lcol.p, bad.p);
      DLAUNCH(k_expand_rows_d, in.nr, in.nr, in.ptr.p, P.halo.send_idx.p, rowkey.p);
      R.offd.nr = nc;
      R.offd.nc = nhr;
      R.offd.nnz = in.nnz;
      R.offd.ptr.alloc((size_t)nc + 1);
      R.offd.col.alloc((size_t)std::max<i64>(in.nnz, 1));
      R.offd.val.alloc((size_t)std::max<i64>(in.nnz, 1));
      DVec<int> cnt((size_t)nc + 1);
      cnt.zero();
      DLAUNCH(k_hist_d, in.nnz, in.nnz, rowkey.p, cnt.p);
      exscan(cnt.p, R.offd.ptr.p, (i64)nc + 1);
      if (in.nnz > 0) {
        DVec<int> iota((size_t)in.nnz), perm((size_t)in.nnz), keys_out((size_t)in.nnz);
        DLAUNCH(k_iota_d, in.nnz, in.nnz, iota.p);
        size_t tb = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, tb, rowkey.p, keys_out.p, iota.p, perm.p, (int)in.nnz, 0, 32, stream());
        DVec<char> tmp(tb);
        cub::DeviceRadixSort::SortPairs(tmp.p, tb, rowkey.p, keys_out.p, iota.p, perm.p, (int)in.nnz, 0, 32, stream());
        count_launch(2);
        // lcol holds merged ids (n + k for halo fine points): store the offd-local index k
        DLAUNCH(k_gather_entries, in.nnz, in.nnz, perm.p, lcol.p, in.val.p, R.offd.col.p, R.offd.val.p);
        DLAUNCH(k_shift_ptr, in.nnz, (int)in.nnz, R.offd.col.p, -n, R.offd.col.p);
      }
      B200_REQUIRE(fetch1(bad.p) == 0, "restriction: received a fine id outside the expected set");
      R.n_halo = nhr;
      R.col_map_offd.resize(nhr);
      if (nhr) {
        B200_CUDA(cudaMemcpyAsync(R.col_map_offd.data(), rmap.p, (size_t)nhr * sizeof(i64), cudaMemcpyDeviceToHost, stream()));
        B200_CUDA(cudaStreamSynchronize(stream()));
      }
    }
    set_par_shape(R, nc, n, cstart, Ap.row_start, gnc, gn);
    build_offd_rows(R);
    build_partition_and_halo(R);

    // Q = A_merged * P_ext
    sub.reset();
    sub.reset(new PhaseTimerD(amg, "rap.AP"));
    DCsr Q;
    DVec<i64> xq;
    int nxq = 0;
    {
      Numbering Np{cstart, nc, xc0.p, nxc0, nullptr, 0};
      RowsMsg extP;
      fetch_halo_rows(Ap, Pm.ptr.p, Pm.col.p, Pm.val.p, nullptr, Np, extP);
      DVec<i64> all((size_t)std::max<i64>(nxc0 + extP.nnz, 1));
      dcopy(all.p, xc0.p, nxc0);
      // own-range ids in the fetched rows are filtered out by marking them as duplicates of an existing entry
      DVec<i64> unk;
      int nunk = collect_unknown(extP.col.p, extP.nnz, Np, unk);
      dcopy(all.p + nxc0, unk.p, nunk);
      nxq = sort_unique(all.p, (i64)nxc0 + nunk, xq);
      Numbering Nq{cstart, nc, xq.p, nxq, nullptr, 0};
      DVec<int> bad(1);
      bad.zero();
      DLAUNCH(k_remap_cols, Pm.nnz, Pm.nnz, Pm.col.p, nc, xc0.p, Nq, bad.p);
      B200_REQUIRE(fetch1(bad.p) == 0, "A*P: column renumbering failed");
      DCsr Pext;
      stack_rows(n, Pm.ptr.p, Pm.col.p, Pm.val.p, Pm.nnz, extP, Nq, nc + nxq, true, Pext);
      PhaseTimerD t2(amg, "rap.AP.spgemm");
      csr_spgemm(Am, Pext, Q, false, false);
    }
    // A_c = R_merged * Q_ext
    sub.reset();
    sub.reset(new PhaseTimerD(amg, "rap.RQ"));
    {
      Numbering Nq{cstart, nc, xq.p, nxq, nullptr, 0};
      RowsMsg extQ;
      fetch_halo_rows(R, Q.ptr.p, Q.col.p, Q.val.p, nullptr, Nq, extQ);
      DVec<i64> unk, xa;
      int nunk = collect_unknown(extQ.col.p, extQ.nnz, Nq, unk);
      DVec<i64> all((size_t)std::max(nxq + nunk, 1));
      dcopy(all.p, xq.p, nxq);
      dcopy(all.p + nxq, unk.p, nunk);
      int nxa = sort_unique(all.p, (i64)nxq + nunk, xa);
      Numbering Na{cstart, nc, xa.p, nxa, nullptr, 0};
      DVec<int> bad(1);
      bad.zero();
      DLAUNCH(k_remap_cols, Q.nnz, Q.nnz, Q.col.p, nc, xq.p, Na, bad.p);
      B200_REQUIRE(fetch1(bad.p) == 0, "R*Q: column renumbering failed");
      DCsr Qext, Rm, Acm;
      {
        PhaseTimerD t2(amg, "rap.RQ.stack");
        stack_rows(n, Q.ptr.p, Q.col.p, Q.val.p, Q.nnz, extQ, Na, nc + nxa, true, Qext);
        merge_parcsr(R, Rm);
      }
      {
        PhaseTimerD t2(amg, "rap.RQ.spgemm");
        csr_spgemm(Rm, Qext, Acm, true);
      }
      Lc.A_own.reset(new ParCsr());
      Lc.A = Lc.A_own.get();
      {
        PhaseTimerD t3(amg, "rap.RQ.split");
        split_merged(Acm, nc, xa.p, nxa, *Lc.A);
      }
      set_par_shape(*Lc.A, nc, nc, cstart, cstart, gnc, gnc);
      PhaseTimerD t4(amg, "rap.RQ.halo_plan");
      build_partition_and_halo(*Lc.A);
    }
  }
  return true;
}

}  // namespace b200
