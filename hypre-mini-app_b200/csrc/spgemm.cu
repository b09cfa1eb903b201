// spgemm.cu -- CSR transposition and the two-pass hash SpGEMM behind the Galerkin product A_c = P^T (A P).
// Replaces hypre's RAP / ParCSR mat-mat kernels reached from HYPRE_BoomerAMGSetup with rap2 = 1
// (/root/reference/src/HypreSystem.cpp:185-188 option, :323 via :692 call); definition: SURVEY.md A.6.
//
// csr_spgemm is Gustavson's algorithm with a fixed floating-point order.  A group of G threads owns one output row;
// the loop over the entries (i,k) of the A row is sequential, the lanes run across the entries (k,j) of row k of B and
// each product is added to the hash slot of column j.  Columns inside one B row are distinct, so no two lanes touch
// the same slot in one k step and every output entry accumulates its terms in exactly the order of the sequential
// algorithm: bit-identical to the CPU oracle, no floating-point atomics.
//   pass 1 (symbolic): rows binned by an upper bound of their product count, distinct columns counted in a
//                      shared-memory hash table sized 2 x bin limit;
//   pass 2 (numeric):  rows binned by their exact length, accumulation as above, compaction, and the row is written in
//                      canonical ParCSR order (diagonal first for square operators, then ascending column).
// Table sizes 64 ... 8192 (numeric) / 32768 (symbolic) slots live in shared memory; G is chosen from the mean row
// length of B (4 ... 32 lanes, a whole block for the largest tables) and the number of rows per block from the table
// footprint so that several blocks stay resident per SM -- the row loop is a chain of dependent loads, and occupancy is
// what hides it.  Rows beyond the largest shared table use global-memory tables.
// Compiled with -fmad=false; the products use explicit __dmul_rn / __dadd_rn anyway.
#include "amg.h"
#include "amg_rows.cuh"
#include <cub/cub.cuh>
#include <algorithm>

namespace b200 {

namespace {

inline rows::Csr rv(const DCsr &A) { return rows::Csr{A.nr, A.nc, A.ptr.p, A.col.p, A.val.p}; }

// ---- transposition: stable sort of the entry indices by column (CUB radix sort) ---------------------------------------
__global__ void k_expand_rows(int nr, const int *ptr, int *rowidx) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nr) return;
  for (int k = ptr[i]; k < ptr[i + 1]; ++k) rowidx[k] = i;
}
__global__ void k_iota(int *a, i64 n) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) a[k] = (int)k;
}
__global__ void k_transpose_gather(i64 nnz, const int *perm, const int *rowidx, const double *val, int *tcol, double *tval) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nnz) return;
  int src = perm[k];
  tcol[k] = rowidx[src];
  if (val) tval[k] = val[src];
}
__global__ void k_hist(const int *col, i64 nnz, int *cnt) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nnz) atomicAdd(cnt + col[k], 1);
}

// ---- SpGEMM -------------------------------------------------------------------------------------------------------------
template <int G>
__device__ __forceinline__ void gsync(unsigned mask) {
  if (G <= 32) __syncwarp(mask);
  else __syncthreads();
}
__device__ __forceinline__ int hash_insert(int *keys, int H, int c, bool &isnew) {
  unsigned h = (((unsigned)c * 2654435761u) >> 11) & (unsigned)(H - 1);
  while (true) {
    int old = atomicCAS(keys + h, -1, c);
    if (old == -1) { isnew = true; return (int)h; }
    if (old == c) { isnew = false; return (int)h; }
    h = (h + 1) & (unsigned)(H - 1);
  }
}

// symbolic pass, optimistic tier: give up after a bounded number of probes (the table is smaller than the row's upper bound)
__device__ __forceinline__ bool hash_insert_capped(int *keys, int H, int c, bool &isnew) {
  unsigned h = (((unsigned)c * 2654435761u) >> 11) & (unsigned)(H - 1);
  for (int probes = 0; probes < 48; ++probes) {
    int old = atomicCAS(keys + h, -1, c);
    if (old == -1) { isnew = true; return true; }
    if (old == c) { isnew = false; return true; }
    h = (h + 1) & (unsigned)(H - 1);
  }
  isnew = false;
  return false;
}

// Rows whose product-count bound is far above their true length (R * Q on the coarser levels: bound ~ 2000, length ~ 70)
// are first counted by one warp in a 256- or 1024-slot table; a row whose table fills up (an insertion gives up) is marked -1 and
// counted again by the exact tier for its bound.  A count that completes is exact, whatever the load of the table.
template <int H_TRY>
__global__ void __launch_bounds__(256) k_spgemm_count_try(rows::Csr A, rows::Csr B, const int *row_list, int nrows, int *nnz_out) {
  __shared__ int s_keys[8][H_TRY];
  const int g = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int idx = blockIdx.x * 8 + g;
  if (idx >= nrows) return;
  const int row = row_list[idx];
  int *keys = s_keys[g];
  for (int s = lane; s < H_TRY; s += 32) keys[s] = -1;
  __syncwarp();
  const int ab = A.ptr[row], ae = A.ptr[row + 1];
  int mine = 0;
  bool fail = false;
  for (int base = ab; base < ae && !fail; base += 32) {
    const int ka = base + lane;
    int bb_l = 0, be_l = 0;
    if (ka < ae) { const int k = A.col[ka]; bb_l = B.ptr[k]; be_l = B.ptr[k + 1]; }
    const int nstep = min(32, ae - base);
    int bb = __shfl_sync(0xffffffffu, bb_l, 0), be = __shfl_sync(0xffffffffu, be_l, 0);
    int c_cur = (bb + lane < be) ? B.col[bb + lane] : -1;
    for (int t = 0; t < nstep; ++t) {
      int bbn = 0, ben = 0, c_nxt = -1;
      if (t + 1 < nstep) {
        bbn = __shfl_sync(0xffffffffu, bb_l, t + 1);
        ben = __shfl_sync(0xffffffffu, be_l, t + 1);
        if (bbn + lane < ben) c_nxt = B.col[bbn + lane];
      }
      bool bad = false;
      if (c_cur >= 0) {
        bool isnew;
        bad |= !hash_insert_capped(keys, H_TRY, c_cur, isnew);
        mine += isnew ? 1 : 0;
      }
      for (int kb = bb + 32 + lane; kb < be; kb += 32) {
        bool isnew;
        bad |= !hash_insert_capped(keys, H_TRY, B.col[kb], isnew);
        mine += isnew ? 1 : 0;
      }
      if (__any_sync(0xffffffffu, bad)) { fail = true; break; }
      bb = bbn; be = ben; c_cur = c_nxt;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mine += __shfl_down_sync(0xffffffffu, mine, o);
  if (lane == 0) nnz_out[row] = fail ? -1 : mine;
}
// size[i] = bound[i] for the rows the optimistic tier gave up on, 0 elsewhere
__global__ void k_retry_size(int n, const int *nnz_row, const int *bound, int *size) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) size[i] = (nnz_row[i] == -1) ? bound[i] : 0;
}

// one output row handled by the G threads of a group.  keys/vals: hash table of H slots; lk/lv: compaction list.
template <int G, bool NUMERIC>
__device__ __forceinline__ void spgemm_row(const rows::Csr &A, const rows::Csr &B, int row, int lane, unsigned mask, int *keys,
                                           double *vals, int H, int *cnt, int *lk, double *lv, const int *Cptr, int *Ccol,
                                           double *Cval, int *nnz_out, int diag_col, bool sorted) {
  // row < 0: a group without a row (tail of the list).  It follows the warp's control flow with an empty row, because for
  // G <= 32 every shuffle / warp barrier below names the FULL warp: a group-specific mask would make the compiler
  // serialise each of them over the groups of the warp (match + per-group barrier), several times the cost of the loop.
  const int ab = row >= 0 ? A.ptr[row] : 0, ae = row >= 0 ? A.ptr[row + 1] : 0;   // requested before the table is cleared
  for (int s = lane; s < H; s += G) {
    keys[s] = -1;
    if (NUMERIC) vals[s] = 0.0;
  }
  if (lane == 0) *cnt = 0;
  gsync<G>(mask);
  int mine = 0;
  if constexpr (G <= 32) {
    // The entries of the A row are loaded G at a time, one per lane, together with the extent of the B row each of them
    // selects.  The k loop itself stays sequential -- it fixes the floating-point order -- but gets its operands by shuffle.
    // Two chunks are in flight: while chunk c is accumulated, the B-row extents of chunk c+1 and the A entries of chunk
    // c+2 are on their way; every step requests the first G entries of the next B row before it accumulates its own.
    // Chunk and step counts are the maxima over the groups of the warp (groups that are done run predicated off).
    constexpr unsigned FULL = 0xffffffffu;
    int k_l = 0, bb_l = 0, be_l = 0;
    double a_l = 0.0;
    if (ab + lane < ae) {
      k_l = A.col[ab + lane];
      if (NUMERIC) a_l = A.val[ab + lane];
      bb_l = B.ptr[k_l];
      be_l = B.ptr[k_l + 1];
    }
    int k_n = 0;
    double a_n = 0.0;
    if (ab + G + lane < ae) { k_n = A.col[ab + G + lane]; if (NUMERIC) a_n = A.val[ab + G + lane]; }
    int bb = __shfl_sync(FULL, bb_l, 0, G), be = __shfl_sync(FULL, be_l, 0, G);
    int c_cur = -1;
    double v_cur = 0.0;
    if (bb + lane < be) { c_cur = B.col[bb + lane]; if (NUMERIC) v_cur = B.val[bb + lane]; }
    const int max_chunks = __reduce_max_sync(FULL, (ae - ab + G - 1) / G);
    for (int c = 0; c < max_chunks; ++c) {
      const int base = ab + c * G;   // >= ae once this group is done
      int bb_n = 0, be_n = 0, k_nn = 0;
      double a_nn = 0.0;
      if (base + G + lane < ae) { bb_n = B.ptr[k_n]; be_n = B.ptr[k_n + 1]; }
      if (base + 2 * G + lane < ae) { k_nn = A.col[base + 2 * G + lane]; if (NUMERIC) a_nn = A.val[base + 2 * G + lane]; }
      const int nstep = max(0, min(G, ae - base));
      const int tmax = __reduce_max_sync(FULL, nstep);
      for (int t = 0; t < tmax; ++t) {
        const double a = NUMERIC ? __shfl_sync(FULL, a_l, t, G) : 0.0;
        int bbn = __shfl_sync(FULL, bb_l, (t + 1) & (G - 1), G), ben = __shfl_sync(FULL, be_l, (t + 1) & (G - 1), G);
        if (t + 1 >= nstep) { bbn = 0; ben = 0; }
        if (t == tmax - 1) {   // warp-uniform: the first B row of the next chunk
          const int xb = __shfl_sync(FULL, bb_n, 0, G), xe = __shfl_sync(FULL, be_n, 0, G);
          if (t + 1 >= nstep && base + G < ae) { bbn = xb; ben = xe; }
        }
        int c_nxt = -1;
        double v_nxt = 0.0;
        if (bbn + lane < ben) { c_nxt = B.col[bbn + lane]; if (NUMERIC) v_nxt = B.val[bbn + lane]; }
        if (t < nstep) {
          if (c_cur >= 0) {
            bool isnew;
            int slot = hash_insert(keys, H, c_cur, isnew);
            if (NUMERIC) vals[slot] = __dadd_rn(vals[slot], __dmul_rn(a, v_cur));
            else mine += isnew ? 1 : 0;
          }
          for (int kb = bb + G + lane; kb < be; kb += G) {  // B rows longer than the group
            bool isnew;
            int slot = hash_insert(keys, H, B.col[kb], isnew);
            if (NUMERIC) vals[slot] = __dadd_rn(vals[slot], __dmul_rn(a, B.val[kb]));
            else mine += isnew ? 1 : 0;
          }
        }
        if (NUMERIC) __syncwarp(FULL);
        bb = bbn; be = ben; c_cur = c_nxt; v_cur = v_nxt;
      }
      a_l = a_n; bb_l = bb_n; be_l = be_n; k_n = k_nn; a_n = a_nn;
    }
  } else {
  for (int ka = ab; ka < ae; ++ka) {
    const int k = A.col[ka];
    const double a = NUMERIC ? A.val[ka] : 0.0;
    const int bb = B.ptr[k], be = B.ptr[k + 1];
    for (int kb = bb + lane; kb < be; kb += G) {
      bool isnew;
      int slot = hash_insert(keys, H, B.col[kb], isnew);
      if (NUMERIC) {
        double prod = __dmul_rn(a, B.val[kb]);
        vals[slot] = __dadd_rn(vals[slot], prod);
      } else {
        mine += isnew ? 1 : 0;
      }
    }
    if (NUMERIC) gsync<G>(mask);
  }
  }
  if constexpr (!NUMERIC) {
    if (G <= 32) {
#pragma unroll
      for (int o = G >> 1; o > 0; o >>= 1) mine += __shfl_down_sync(mask, mine, o, G);
      if (lane == 0 && row >= 0) nnz_out[row] = mine;
    } else {
      if (mine) atomicAdd(cnt, mine);
      __syncthreads();
      if (lane == 0) nnz_out[row] = *cnt;
      __syncthreads();
    }
    return;
  } else {
  const int base = row >= 0 ? Cptr[row] : 0;
  if (!sorted) {
    // intermediate product (Q = A P): the consumer R * Q accumulates over the entries of R's rows, at most one term per
    // Q row for any output entry, so the order inside a Q row is irrelevant and the row is written as it is compacted
    for (int s = lane; s < H; s += G) {
      int c = keys[s];
      if (c != -1) {
        int p = atomicAdd(cnt, 1);
        Ccol[base + p] = c;
        Cval[base + p] = vals[s];
      }
    }
    gsync<G>(mask);
    return;
  }
  // compaction; the list holds sort keys (the diagonal column becomes -1 so that it ranks first)
  for (int s = lane; s < H; s += G) {
    int c = keys[s];
    if (c != -1) {
      int p = atomicAdd(cnt, 1);
      lk[p] = (c == diag_col) ? -1 : c;
      lv[p] = vals[s];
    }
  }
  gsync<G>(mask);
  const int n = *cnt;
  if (lane == 0 && (n & 1)) lk[n] = 0x7fffffff;   // pad to an even length (capacity H / 2 is even, n odd < H / 2)
  gsync<G>(mask);
  // rank = number of smaller keys (keys are distinct); two keys per shared-memory load
  for (int t = lane; t < n; t += G) {
    const int key = lk[t];
    int rank = 0;
#pragma unroll 4
    for (int u = 0; u < n; u += 2) {
      const int2 kk = *reinterpret_cast<const int2 *>(lk + u);
      rank += (kk.x < key) ? 1 : 0;
      rank += (kk.y < key) ? 1 : 0;
    }
    Ccol[base + rank] = (key == -1) ? diag_col : key;
    Cval[base + rank] = lv[t];
  }
  gsync<G>(mask);
  }
}

// shared-memory table kernels.  G threads per row, blockDim.x / G rows per block.
template <int G, bool NUMERIC>
__global__ void k_spgemm_smem(rows::Csr A, rows::Csr B, const int *row_list, int nrows, int H, const int *Cptr, int *Ccol,
                              double *Cval, int *nnz_out, int diag_first, int sorted) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int groups = blockDim.x / G;
  const int g = threadIdx.x / G, lane = threadIdx.x % G;
  // per group: cnt (int, padded to 8 B) | [vals H doubles | lv H/2 doubles] | keys H ints | [lk H/2 ints]; the compaction
  // list lv / lk exists only for sorted output (an unsorted row is written straight from the table)
  const bool list = NUMERIC && sorted != 0;
  size_t per = 8 + (NUMERIC ? ((size_t)H * 12 + (list ? (size_t)H * 6 : 0)) : (size_t)H * 4);
  unsigned char *base = smem + per * g;
  int *cnt = (int *)base;
  double *vals = (double *)(base + 8);
  double *lv = vals + (NUMERIC ? H : 0);
  int *keys = (int *)(base + 8 + (NUMERIC ? (size_t)H * (list ? 12 : 8) : 0));
  int *lk = keys + H;
  const unsigned mask = 0xffffffffu;   // G <= 32: all groups of a warp move together (see spgemm_row)
  int idx = blockIdx.x * groups + g;
  if (G > 32 && idx >= nrows) return;  // groups == 1, uniform for the whole block
  int row = idx < nrows ? row_list[idx] : -1;
  spgemm_row<G, NUMERIC>(A, B, row, lane, mask, keys, vals, H, cnt, lk, lv, Cptr, Ccol, Cval, nnz_out, diag_first ? row : -2, sorted != 0);
}

// global-memory table kernel: one table per block, grid-stride over the bin
template <bool NUMERIC>
__global__ void k_spgemm_gmem(rows::Csr A, rows::Csr B, const int *row_list, int nrows, int H, int *gkeys, double *gvals, int *glk,
                              double *glv, const int *Cptr, int *Ccol, double *Cval, int *nnz_out, int diag_first, int sorted) {
  __shared__ int cnt;
  int *keys = gkeys + (size_t)blockIdx.x * H;
  double *vals = NUMERIC ? gvals + (size_t)blockIdx.x * H : nullptr;
  int *lk = NUMERIC ? glk + (size_t)blockIdx.x * H : nullptr;
  double *lv = NUMERIC ? glv + (size_t)blockIdx.x * H : nullptr;
  for (int idx = blockIdx.x; idx < nrows; idx += gridDim.x) {
    int row = row_list[idx];
    spgemm_row<256, NUMERIC>(A, B, row, threadIdx.x, 0, keys, vals, H, &cnt, lk, lv, Cptr, Ccol, Cval, nnz_out,
                              diag_first ? row : -2, sorted != 0);
  }
}
__global__ void k_spgemm_bound(rows::Csr A, rows::Csr B, int *bound) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.nr) return;
  i64 s = 0;
  for (int k = A.ptr[i]; k < A.ptr[i + 1]; ++k) {
    int c = A.col[k];
    s += B.ptr[c + 1] - B.ptr[c];
  }
  if (s > B.nc) s = B.nc;
  bound[i] = (int)s;
}

// ---- row binning: bin b holds the rows with lim[b-1] < size <= lim[b]; rows of size 0 go nowhere; bin NB = the rest ------
constexpr int NB = 7;
struct BinLimits { int lim[NB]; };
__device__ __forceinline__ int bin_of(int s, const BinLimits &L) {
  int b = NB;
#pragma unroll
  for (int k = NB - 1; k >= 0; --k)
    if (s <= L.lim[k]) b = k;
  return b;
}
// d[0..NB]: counts, d[8..8+NB]: offsets, d[16..16+NB]: cursors, d[24]: max size in the last bin
__global__ void k_bin_count(int n, const int *size, BinLimits L, int *d) {
  __shared__ int sc[NB + 1];
  if (threadIdx.x <= NB) sc[threadIdx.x] = 0;
  __syncthreads();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    int s = size[i];
    if (s > 0) atomicAdd(&sc[bin_of(s, L)], 1);
  }
  __syncthreads();
  if (threadIdx.x <= NB && sc[threadIdx.x]) atomicAdd(d + threadIdx.x, sc[threadIdx.x]);
}
__global__ void k_bin_offsets(int *d) {
  int off = 0;
  for (int k = 0; k <= NB; ++k) { d[8 + k] = off; off += d[k]; }
}
__global__ void k_bin_fill(int n, const int *size, BinLimits L, int *d, int *list) {
  __shared__ int sc[NB + 1], sbase[NB + 1], smax;
  if (threadIdx.x <= NB) sc[threadIdx.x] = 0;
  if (threadIdx.x == 0) smax = 0;
  __syncthreads();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int b = -1, local = 0;
  if (i < n) {
    int s = size[i];
    if (s > 0) {
      b = bin_of(s, L);
      local = atomicAdd(&sc[b], 1);
      if (b == NB) atomicMax(&smax, s);
    }
  }
  __syncthreads();
  if (threadIdx.x <= NB && sc[threadIdx.x]) sbase[threadIdx.x] = d[8 + threadIdx.x] + atomicAdd(d + 16 + threadIdx.x, sc[threadIdx.x]);
  if (threadIdx.x == 0 && smax) atomicMax(d + 24, smax);
  __syncthreads();
  if (b >= 0) list[sbase[b] + local] = i;
}

struct Bins {
  DVec<int> list;
  int count[NB + 1], offset[NB + 1];
  int max_size = 0;
};
void make_bins(int n, const int *size, const BinLimits &L, Bins &b) {
  DVec<int> d(32);
  d.zero();
  b.list.alloc((size_t)std::max(n, 1));
  k_bin_count<<<div_up(n, 256), 256, 0, stream()>>>(n, size, L, d.p);
  k_bin_offsets<<<1, 1, 0, stream()>>>(d.p);
  k_bin_fill<<<div_up(n, 256), 256, 0, stream()>>>(n, size, L, d.p, b.list.p);
  count_launch(3);
  std::vector<int> h = d.to_host();
  for (int k = 0; k <= NB; ++k) { b.count[k] = h[k]; b.offset[k] = h[8 + k]; }
  b.max_size = h[24];
}

inline int next_pow2(i64 v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

template <int G, bool NUMERIC>
void launch_smem(const rows::Csr &a, const rows::Csr &b, const int *list, int count, int H, int gpb, size_t smem, const int *Cptr,
                 int *Ccol, double *Cval, int *nnz_out, bool diag_first, bool sorted) {
  auto kern = k_spgemm_smem<G, NUMERIC>;
  static bool attr_set = false;  // per instantiation
  if (!attr_set) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set = true;
  }
  kern<<<div_up(count, gpb), gpb * G, smem, stream()>>>(a, b, list, count, H, Cptr, Ccol, Cval, nnz_out, diag_first ? 1 : 0, sorted ? 1 : 0);
  count_launch();
}

template <bool NUMERIC>
void spgemm_pass(const DCsr &A, const DCsr &B, const Bins &bins, const BinLimits &L, const int *Cptr, int *Ccol, double *Cval,
                 int *nnz_out, bool diag_first, bool sorted) {
  rows::Csr a = rv(A), b = rv(B);
  // lanes per row from the mean row length of B
  double mean_b = B.nr > 0 ? (double)B.nnz / B.nr : 1.0;
  int g_b = 4;
  while (g_b < 32 && g_b < mean_b) g_b *= 2;
  for (int k = 0; k < NB; ++k) {
    if (!bins.count[k]) continue;
    const int H = 2 * next_pow2(L.lim[k]);
    const size_t per = 8 + (NUMERIC ? (size_t)H * (sorted ? 18 : 12) : (size_t)H * 4);
    const int *list = bins.list.p + bins.offset[k];
    int G = g_b;
    if (H >= 4096) G = 256;
    else if (H >= 1024) G = 32;
    int gpb = 1;
    if (G <= 32) {
      int maxg = 256 / G;
      size_t budget = 72 * 1024;
      while (gpb * 2 <= maxg && (size_t)(gpb * 2) * per <= budget) gpb *= 2;
      while (gpb * G < 64 && G < 32) G *= 2;  // at least two warps per block
    }
    size_t smem = per * gpb;
    switch (G) {
      case 4: launch_smem<4, NUMERIC>(a, b, list, bins.count[k], H, gpb, smem, Cptr, Ccol, Cval, nnz_out, diag_first, sorted); break;
      case 8: launch_smem<8, NUMERIC>(a, b, list, bins.count[k], H, gpb, smem, Cptr, Ccol, Cval, nnz_out, diag_first, sorted); break;
      case 16: launch_smem<16, NUMERIC>(a, b, list, bins.count[k], H, gpb, smem, Cptr, Ccol, Cval, nnz_out, diag_first, sorted); break;
      case 32: launch_smem<32, NUMERIC>(a, b, list, bins.count[k], H, gpb, smem, Cptr, Ccol, Cval, nnz_out, diag_first, sorted); break;
      default: launch_smem<256, NUMERIC>(a, b, list, bins.count[k], H, 1, per, Cptr, Ccol, Cval, nnz_out, diag_first, sorted); break;
    }
  }
  if (bins.count[NB]) {  // global-memory tables
    int H = next_pow2(2 * (i64)bins.max_size);
    int grid = std::min(bins.count[NB], 2 * ctx().sm_count);
    DVec<int> gk((size_t)grid * H), glk;
    DVec<double> gv, glv;
    if (NUMERIC) { gv.alloc((size_t)grid * H); glk.alloc((size_t)grid * H); glv.alloc((size_t)grid * H); }
    k_spgemm_gmem<NUMERIC><<<grid, 256, 0, stream()>>>(a, b, bins.list.p + bins.offset[NB], bins.count[NB], H, gk.p, gv.p, glk.p,
                                                       glv.p, Cptr, Ccol, Cval, nnz_out, diag_first, sorted ? 1 : 0);
    count_launch();
    B200_CUDA(cudaStreamSynchronize(stream()));
  }
  B200_KERNEL_CHECK();
}

template <class T>
void exclusive_scan(const T *in, T *out, i64 n) {
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, in, out, (int)n, stream());
  DVec<char> tmp(tb);
  cub::DeviceScan::ExclusiveSum(tmp.p, tb, in, out, (int)n, stream());
  count_launch(2);
}

__global__ void k_sum_i64(const int *__restrict__ v, i64 n, unsigned long long *out) {
  unsigned long long s = 0;
  for (i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (i64)gridDim.x * blockDim.x) s += (unsigned long long)v[k];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0 && s) atomicAdd(out, s);
}

}  // namespace

// Sum of n non-negative per-row counts in 64 bits.  Row pointers are int32 (DCsr), so every scan of counts that a product,
// an interpolation or a transposition produces is guarded with this first: a total beyond int32 fails with a diagnostic
// instead of wrapping silently and under-allocating the column / value arrays.
i64 checked_total(const int *counts, i64 n, const char *what) {
  if (n <= 0) return 0;
  DVec<unsigned long long> acc(1);
  acc.zero();
  k_sum_i64<<<std::min<int>(div_up(n, 256), 4 * ctx().sm_count), 256, 0, stream()>>>(counts, n, acc.p);
  count_launch();
  unsigned long long h = 0;
  B200_CUDA(cudaMemcpyAsync(&h, acc.p, sizeof(h), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  B200_REQUIRE(h < 2147483640ull, std::string(what) + ": " + std::to_string(h) +
                                      " entries exceed the 32-bit row pointers of one rank; partition the system over more GPUs");
  return (i64)h;
}

// T = A^T; entries of every T row keep the order of A's rows (stable).  val may be absent (pattern only) when A.val.n == 0.
void csr_transpose(const DCsr &A, DCsr &T) {
  T.nr = A.nc;
  T.nc = A.nr;
  T.nnz = A.nnz;
  T.ptr.alloc((size_t)T.nr + 1);
  T.col.alloc((size_t)std::max<i64>(A.nnz, 1));
  B200_REQUIRE(A.nnz < 2147483640LL, "csr_transpose: more than 2^31 entries on one rank");
  const bool has_val = A.val.n > 0;
  if (has_val) T.val.alloc((size_t)std::max<i64>(A.nnz, 1));
  DVec<int> cnt((size_t)T.nr + 1);
  cnt.zero();
  if (A.nnz == 0) { T.ptr.zero(); return; }
  k_hist<<<div_up(A.nnz, 256), 256, 0, stream()>>>(A.col.p, A.nnz, cnt.p);
  exclusive_scan(cnt.p, T.ptr.p, (i64)T.nr + 1);
  DVec<int> rowidx((size_t)A.nnz), iota((size_t)A.nnz), perm((size_t)A.nnz), keys_out((size_t)A.nnz);
  k_expand_rows<<<div_up(A.nr, 256), 256, 0, stream()>>>(A.nr, A.ptr.p, rowidx.p);
  k_iota<<<div_up(A.nnz, 256), 256, 0, stream()>>>(iota.p, A.nnz);
  int bits = 1;
  while ((1LL << bits) < A.nc) ++bits;
  size_t tb = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tb, A.col.p, keys_out.p, iota.p, perm.p, (int)A.nnz, 0, bits, stream());
  DVec<char> tmp(tb);
  cub::DeviceRadixSort::SortPairs(tmp.p, tb, A.col.p, keys_out.p, iota.p, perm.p, (int)A.nnz, 0, bits, stream());
  k_transpose_gather<<<div_up(A.nnz, 256), 256, 0, stream()>>>(A.nnz, perm.p, rowidx.p, has_val ? A.val.p : nullptr, T.col.p,
                                                                has_val ? T.val.p : nullptr);
  count_launch(6);
  B200_KERNEL_CHECK();
}

void csr_spgemm(const DCsr &A, const DCsr &B, DCsr &C, bool diag_first, bool sorted) {
  B200_REQUIRE(A.nc == B.nr, "csr_spgemm: dimension mismatch");
  C.nr = A.nr;
  C.nc = B.nc;
  C.ptr.alloc((size_t)C.nr + 1);
  if (A.nr == 0) { C.nnz = 0; C.ptr.zero(); return; }
  const double in_bytes = 12.0 * (double)(A.nnz + B.nnz) + 4.0 * (double)(A.nr + B.nr);
  DVec<int> bound((size_t)A.nr + 1), nnz_row((size_t)A.nr + 1);
  nnz_row.zero();
  B200_CUDA(cudaMemsetAsync(bound.p + A.nr, 0, sizeof(int), stream()));
  {
    ProfScope prof("spgemm_symbolic", in_bytes + 4.0 * A.nr);
    k_spgemm_bound<<<div_up(A.nr, 256), 256, 0, stream()>>>(rv(A), rv(B), bound.p);
    count_launch();
    Bins bins;
    BinLimits L{{16, 32, 64, 128, 512, 2048, 16384}};
    make_bins(A.nr, bound.p, L, bins);
    // optimistic warp-per-row count first, exact tier for the failures: bin 4 (128 < bound <= 512) in a 256-slot table,
    // bins 5, 6 and the global-memory bin (bound > 512) in a 1024-slot table
    const int first_big = 5;
    const int n_mid = bins.count[4];
    const int n_big = (bins.offset[NB] + bins.count[NB]) - bins.offset[first_big];
    if (n_mid > 0) {
      k_spgemm_count_try<256><<<div_up(n_mid, 8), 256, 0, stream()>>>(rv(A), rv(B), bins.list.p + bins.offset[4], n_mid, nnz_row.p);
      count_launch();
      bins.count[4] = 0;
    }
    if (n_big > 0) {
      k_spgemm_count_try<1024><<<div_up(n_big, 8), 256, 0, stream()>>>(rv(A), rv(B), bins.list.p + bins.offset[first_big], n_big, nnz_row.p);
      count_launch();
      for (int k = first_big; k <= NB; ++k) bins.count[k] = 0;
    }
    spgemm_pass<false>(A, B, bins, L, nullptr, nullptr, nullptr, nnz_row.p, diag_first, sorted);
    if (n_mid + n_big > 0) {
      DVec<int> retry((size_t)A.nr);
      k_retry_size<<<div_up(A.nr, 256), 256, 0, stream()>>>(A.nr, nnz_row.p, bound.p, retry.p);
      count_launch();
      Bins again;
      make_bins(A.nr, retry.p, L, again);
      spgemm_pass<false>(A, B, again, L, nullptr, nullptr, nullptr, nnz_row.p, diag_first, sorted);
    }
  }
  C.nnz = checked_total(nnz_row.p, C.nr, "csr_spgemm (Galerkin product)");
  exclusive_scan(nnz_row.p, C.ptr.p, (i64)C.nr + 1);
  C.col.alloc((size_t)std::max<i64>(C.nnz, 1));
  C.val.alloc((size_t)std::max<i64>(C.nnz, 1));
  {
    ProfScope prof("spgemm_numeric", in_bytes + 12.0 * (double)C.nnz + 4.0 * C.nr);
    Bins bins;
    BinLimits L{{16, 32, 64, 128, 256, 1024, 4096}};
    make_bins(A.nr, nnz_row.p, L, bins);
    spgemm_pass<true>(A, B, bins, L, C.ptr.p, C.col.p, C.val.p, nullptr, diag_first, sorted);
  }
}

}  // namespace b200
