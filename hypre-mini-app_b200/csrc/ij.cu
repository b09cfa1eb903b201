// ij.cu -- IJ matrix / vector assembly on the device: COO triplets (global ids) -> ParCSR (diag + offd).
// Replaces HYPRE_IJMatrixSetValues2 / AddToValues2 / Assemble and HYPRE_IJVectorSetValues / GetValues as driven by
// /root/reference/src/HypreSystem.cpp:897-1015 (one column per "row", ncols == NULL, host or device pointers),
// :600-636 (assemble) and :1567-1596 (generator: Set for the diag block, AddTo for the offd block).
// Semantics: Set overwrites, AddTo accumulates, in call order; rows sorted by (row, col); diag block
// in local column numbering with the diagonal entry first; offd columns compressed through the sorted col_map_offd.
// Fast path: a single staged chunk that is already strictly sorted by (row, col) skips the radix sorts.
#include "amg.h"
#include <cub/cub.cuh>
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <sched.h>
#include <thread>

namespace b200 {

void ij_add_int(int n, const int *a, const int *b, int *c);
void build_partition_and_halo(ParCsr &A);

namespace {

bool is_device_ptr(const void *p) {
  if (!p) return false;
  cudaPointerAttributes at;
  cudaError_t e = cudaPointerGetAttributes(&at, p);
  if (e != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

std::atomic<unsigned long long> g_h2d_bytes{0};   // bytes copied from caller host buffers to the device by the IJ interface

template <class T>
void stage_copy(DVec<T> &dst, const T *src, i64 n) {
  dst.alloc((size_t)n);
  if (n == 0) return;
  const bool dev = is_device_ptr(src);
  if (!dev) g_h2d_bytes += (unsigned long long)n * sizeof(T);
  B200_CUDA(cudaMemcpyAsync(dst.p, src, n * sizeof(T), dev ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, stream()));
}

// ---- host rows -> row pointer: the row ids of a triplet list that is sorted by row (the mini-app's loaders and generators produce
// that, /root/reference/src/HypreSystem.cpp:897-955) carry nloc + 1 numbers of information in nnz 64-bit words.  When the caller's
// arrays live on the host the link is the bottleneck of SetValues2 (2.9 GB for 256^3 at ~56 GB/s), so the host threads turn `rows`
// into a local row pointer while the DMA engine ships cols and vals, and only the pointer crosses the link; a kernel expands it
// again.  Anything else (unsorted rows, a row outside [ilower, iupper]) takes the plain copy and is judged by Assemble.
struct PinnedInts {
  int *p = nullptr;
  size_t cap = 0;
  int *get(size_t n) {
    if (n > cap) {
      if (p) cudaFreeHost(p);
      p = nullptr;
      cap = 0;
      if (cudaHostAlloc((void **)&p, n * sizeof(int), cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); p = nullptr; return nullptr; }
      cap = n;
    }
    return p;
  }
};
PinnedInts g_ptr_stage;   // grow-only, one per process (one rank per process)

int host_threads_for_staging() {
  cpu_set_t set;
  int avail = (int)std::thread::hardware_concurrency();
  if (sched_getaffinity(0, sizeof(set), &set) == 0) avail = CPU_COUNT(&set);
  int t = avail / std::max(1, ctx().nranks);
  return std::max(1, std::min(t, 16));
}

// ptr[r - ilower] = first position of row r (nloc + 1 entries); false if rows are not nondecreasing or leave [ilower, iupper]
bool host_rows_to_ptr(const i64 *rows, i64 n, i64 ilower, i64 iupper, int *ptr) {
  const int T = host_threads_for_staging();
  std::atomic<int> bad{0};
  auto work = [&](int t) {
    const i64 a = n * t / T, b = n * (t + 1) / T;
    if (a >= b) return;
    i64 prev;
    i64 k = a;
    if (a == 0) {
      prev = rows[0];
      if (prev < ilower || prev > iupper) { bad = 1; return; }
      for (i64 r = ilower; r <= prev; ++r) ptr[r - ilower] = 0;
      k = 1;
    } else {
      prev = rows[a - 1];
    }
    for (; k < b; ++k) {
      const i64 cur = rows[k];
      if (cur != prev) {
        if (cur < prev || cur > iupper || prev < ilower) { bad = 1; return; }
        for (i64 r = prev + 1; r <= cur; ++r) ptr[r - ilower] = (int)k;
        prev = cur;
      }
    }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < T; ++t) th.emplace_back(work, t);
  work(0);
  for (auto &x : th) x.join();
  if (bad.load()) return false;
  const i64 last = rows[n - 1];
  for (i64 r = last + 1; r <= iupper + 1; ++r) ptr[r - ilower] = (int)n;
  return true;
}

__global__ void k_expand_rows(int nloc, const int *__restrict__ ptr, i64 ilower, i64 *rows) {
  // one warp per row: rows are a handful of entries
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= nloc) return;
  const i64 g = ilower + r;
  for (int k = ptr[r] + lane; k < ptr[r + 1]; k += 32) rows[k] = g;
}

// stages `rows` (host pointer) through the row pointer; returns false when the plain copy has to be used instead
bool stage_rows_compressed(DVec<i64> &dst, const i64 *rows, i64 n, i64 ilower, i64 iupper) {
  const i64 nloc = iupper - ilower + 1;
  if (n < (1 << 20) || n >= (1LL << 31) - 8 || nloc <= 0 || nloc >= (1LL << 31) - 8 || nloc > n) return false;
  // the scan has to hide behind the transfer of cols and vals (16 threads: 12 ms per 117 M ids, against 34 ms of DMA): with fewer
  // than 6 threads per rank (many ranks on a small host) the plain copy is the faster way
  if (host_threads_for_staging() < 6) return false;
  int *hptr = g_ptr_stage.get((size_t)nloc + 1);
  if (!hptr) return false;
  const bool timing = getenv("B200_IJ_TIMING") != nullptr;
  const auto t0 = std::chrono::steady_clock::now();
  const bool sorted = host_rows_to_ptr(rows, n, ilower, iupper, hptr);
  if (timing)
    fprintf(stderr, "[b200hypre] IJ staging: %lld row ids -> row pointer on %d host threads in %.2f ms (%s)\n", (long long)n,
            host_threads_for_staging(), std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(),
            sorted ? "sorted" : "not sorted: plain copy");
  if (!sorted) return false;
  DVec<int> dptr((size_t)nloc + 1);
  dst.alloc((size_t)n);
  g_h2d_bytes += (unsigned long long)(nloc + 1) * sizeof(int);
  B200_CUDA(cudaMemcpyAsync(dptr.p, hptr, (size_t)(nloc + 1) * sizeof(int), cudaMemcpyHostToDevice, stream()));
  k_expand_rows<<<div_up((i64)nloc * 32, 256), 256, 0, stream()>>>((int)nloc, dptr.p, ilower, dst.p);
  count_launch();
  // the pinned pointer is reused by the next call: it must have left the host before this function returns
  B200_CUDA(cudaStreamSynchronize(stream()));
  return true;
}

__global__ void k_check_sorted(i64 n, const i64 *rows, const i64 *cols, int *bad) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k == 0 || k >= n) return;
  i64 r0 = rows[k - 1], r1 = rows[k];
  if (r1 < r0 || (r1 == r0 && cols[k] <= cols[k - 1])) *bad = 1;
}
// Row-sorted input whose columns are in arbitrary order inside a row (the generator's stencil order on a multi-rank
// grid, many MatrixMarket files): bad[0] = 1 if the rows are not nondecreasing.
__global__ void k_check_rows_sorted(i64 n, const i64 *rows, int *bad) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k == 0 || k >= n) return;
  if (rows[k] < rows[k - 1]) *bad = 1;
}
// The thread that sees the head of a row run sorts that run by column, in place (stable insertion sort; runs are a few
// dozen entries at most).  Runs longer than `maxlen` are left alone and reported through toolong.
__global__ void k_sort_short_rows(i64 n, const i64 *__restrict__ rows, i64 *cols, double *vals, int maxlen, int *toolong) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const i64 r = rows[k];
  if (k > 0 && rows[k - 1] == r) return;
  i64 e = k + 1;
  while (e < n && e - k <= maxlen && rows[e] == r) ++e;
  if (e - k > maxlen) { *toolong = 1; return; }
  for (i64 a = k + 1; a < e; ++a) {
    i64 c = cols[a];
    double v = vals[a];
    i64 b = a - 1;
    while (b >= k && cols[b] > c) { cols[b + 1] = cols[b]; vals[b + 1] = vals[b]; --b; }
    cols[b + 1] = c;
    vals[b + 1] = v;
  }
}
__global__ void k_check_range(i64 n, const i64 *rows, i64 lo, i64 hi, int *bad) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n && (rows[k] < lo || rows[k] > hi)) *bad = 1;
}
__global__ void k_pack_keys(i64 n, const i64 *rows, const i64 *cols, i64 row0, int cbits, i64 *key) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) key[k] = ((rows[k] - row0) << cbits) | cols[k];
}
__global__ void k_iota64(i64 n, int *a) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) a[k] = (int)k;
}
template <class T>
__global__ void k_gather(i64 n, const int *perm, const T *in, T *out) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) out[k] = in[perm[k]];
}
// after the stable sort by (row, col): head[k] = 1 where a new (row, col) key starts
__global__ void k_heads(i64 n, const i64 *rows, const i64 *cols, int *head) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  head[k] = (k == 0 || rows[k] != rows[k - 1] || cols[k] != cols[k - 1]) ? 1 : 0;
}
// combine duplicates in insertion order: set -> v, add -> += v.  One thread per unique key (slot = scan(head)).
__global__ void k_combine(i64 n, const int *head, const int *slot, const i64 *rows, const i64 *cols, const double *vals,
                          const signed char *ops, i64 *urows, i64 *ucols, double *uvals) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n || !head[k]) return;
  double v = 0.0;
  i64 j = k;
  do {
    v = ops[j] ? (v + vals[j]) : vals[j];
    ++j;
  } while (j < n && !head[j]);
  int s = slot[k];
  urows[s] = rows[k];
  ucols[s] = cols[k];
  uvals[s] = v;
}
// per-row counts of diag / offd entries (rows are global ids, sorted)
__global__ void k_row_counts(i64 n, const i64 *rows, const i64 *cols, i64 ilower, i64 jlower, i64 jupper, int *dcnt, int *ocnt) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  int r = (int)(rows[k] - ilower);
  i64 c = cols[k];
  if (c >= jlower && c <= jupper) atomicAdd(dcnt + r, 1);
  else atomicAdd(ocnt + r, 1);
}
__device__ __forceinline__ int lower_bound_i64(const i64 *a, int n, i64 key) {
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = lo + ((hi - lo) >> 1);
    if (a[mid] < key) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}
// thread per row: split the sorted row into diag (diagonal first) and offd (compressed columns)
__global__ void k_split_fill(int nrows, const int *rowstart, const i64 *cols, const double *vals, i64 ilower, i64 jlower, i64 jupper,
                             const int *dptr, int *dcol, double *dval, const int *optr, int *ocol, double *oval,
                             const i64 *colmap, int ncolmap) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nrows) return;
  int b = rowstart[r], e = rowstart[r + 1];
  i64 gdiag = ilower + r;
  int dp = dptr[r], op = optr ? optr[r] : 0;
  // diagonal first
  int dk = -1;
  for (int k = b; k < e; ++k)
    if (cols[k] == gdiag) { dk = k; break; }
  if (dk >= 0 && gdiag >= jlower && gdiag <= jupper) {
    dcol[dp] = (int)(gdiag - jlower);
    dval[dp] = vals[dk];
    ++dp;
  } else dk = -1;
  for (int k = b; k < e; ++k) {
    if (k == dk) continue;
    i64 c = cols[k];
    if (c >= jlower && c <= jupper) {
      dcol[dp] = (int)(c - jlower);
      dval[dp] = vals[k];
      ++dp;
    } else {
      ocol[op] = lower_bound_i64(colmap, ncolmap, c);
      oval[op] = vals[k];
      ++op;
    }
  }
}
__global__ void k_collect_offd(i64 n, const i64 *cols, i64 jlower, i64 jupper, i64 *out, int *counter) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  i64 c = cols[k];
  if (c < jlower || c > jupper) out[atomicAdd(counter, 1)] = c;
}
__global__ void k_offd_rows(int n, const int *optr, int *flag) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flag[i] = optr[i + 1] > optr[i] ? 1 : 0;
  if (i == n) flag[i] = 0;
}
__global__ void k_compact_rows(int n, const int *flag, const int *pos, int *list) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && flag[i]) list[pos[i]] = i;
}

template <class T>
void exscan(const T *in, T *out, i64 n) {
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, in, out, (int)n, stream());
  DVec<char> tmp(tb);
  cub::DeviceScan::ExclusiveSum(tmp.p, tb, in, out, (int)n, stream());
  count_launch(2);
}
template <class T>
T fetch1(const T *dev) {
  T h;
  B200_CUDA(cudaMemcpyAsync(&h, dev, sizeof(T), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  return h;
}

__global__ void k_vec_set(i64 n, const i64 *idx, const double *vals, i64 lo, i64 hi, double *v, int op, i64 seq_start) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  i64 g = idx ? idx[k] : (seq_start + k);
  if (g < lo || g > hi) return;
  if (op) v[g - lo] += vals[k];
  else v[g - lo] = vals[k];
}
__global__ void k_vec_get(i64 n, const i64 *idx, double *out, i64 lo, i64 hi, const double *v) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  i64 g = idx ? idx[k] : (lo + k);
  out[k] = (g >= lo && g <= hi) ? v[g - lo] : 0.0;
}

}  // namespace

void ij_matrix_set(IJMatrix &M, i64 nrows, const int *ncols, const i64 *rows, const int *row_indexes, const i64 *cols,
                   const double *vals, int op) {
  ensure_init();
  if (nrows <= 0) return;
  IJMatrix::Chunk ch;
  ch.op = op;
  if (ncols == nullptr) {  // one entry per "row" (the mini-app's usage)
    ch.n = nrows;
    // cols and vals first: their transfers run while the host threads look at the rows
    stage_copy(ch.cols, cols, nrows);
    stage_copy(ch.vals, vals, nrows);
    if (is_device_ptr(rows) || !stage_rows_compressed(ch.rows, rows, nrows, M.ilower, M.iupper)) stage_copy(ch.rows, rows, nrows);
  } else {  // general CSR-like call: expand on the host (host pointers only)
    B200_REQUIRE(!is_device_ptr(ncols) && !is_device_ptr(rows) && !is_device_ptr(cols) && !is_device_ptr(vals),
                 "IJMatrixSetValues2 with ncols != NULL expects host pointers");
    std::vector<i64> r, c;
    std::vector<double> v;
    i64 off = 0;
    for (i64 i = 0; i < nrows; ++i) {
      i64 start = row_indexes ? row_indexes[i] : off;
      for (int k = 0; k < ncols[i]; ++k) { r.push_back(rows[i]); c.push_back(cols[start + k]); v.push_back(vals[start + k]); }
      off += ncols[i];
    }
    ch.n = (i64)r.size();
    ch.rows.from_host(r);
    ch.cols.from_host(c);
    ch.vals.from_host(v);
    B200_CUDA(cudaStreamSynchronize(stream()));
  }
  M.assembled = false;
  M.staged.push_back(std::move(ch));
}

unsigned long long ij_h2d_bytes() { return g_h2d_bytes.load(); }

void ij_matrix_assemble(IJMatrix &M) {
  ensure_init();
  if (M.assembled && M.staged.empty()) return;
  const int nranks = ctx().nranks;
  i64 ntot = 0;
  for (auto &c : M.staged) ntot += c.n;
  B200_REQUIRE(ntot < (1LL << 31) - 8, "IJ assemble: more than 2^31 staged entries on one rank");
  const int nloc = (int)(M.iupper - M.ilower + 1);
  ParCsr &A = M.par;
  A.n = nloc;
  A.n_cols_local = (int)(M.jupper - M.jlower + 1);
  A.row_start = M.ilower;
  A.col_start = M.jlower;
  // sorted, unique COO
  DVec<i64> rows, cols;
  DVec<double> vals;
  i64 nu = 0;
  bool fast = false;
  DVec<int> bad(1);
  if (M.staged.size() == 1 && M.staged[0].op == 0) {
    auto &c = M.staged[0];
    bad.zero();
    if (c.n) k_check_sorted<<<div_up(c.n, 256), 256, 0, stream()>>>(c.n, c.rows.p, c.cols.p, bad.p);
    count_launch();
    fast = (fetch1(bad.p) == 0);
    if (!fast && c.n) {  // rows in order, columns not: sort every (short) row in place instead of radix-sorting all keys
      DVec<int> flags(2);
      flags.zero();
      k_check_rows_sorted<<<div_up(c.n, 256), 256, 0, stream()>>>(c.n, c.rows.p, flags.p);
      count_launch();
      if (fetch1(flags.p) == 0) {
        k_sort_short_rows<<<div_up(c.n, 256), 256, 0, stream()>>>(c.n, c.rows.p, c.cols.p, c.vals.p, 64, flags.p + 1);
        bad.zero();
        k_check_sorted<<<div_up(c.n, 256), 256, 0, stream()>>>(c.n, c.rows.p, c.cols.p, bad.p);
        count_launch(2);
        // duplicates or a long row fall through to the general path (the in-place sort was stable, so Set/Add order holds)
        fast = (fetch1(flags.p + 1) == 0) && (fetch1(bad.p) == 0);
      }
    }
  }
  if (fast) {
    rows = std::move(M.staged[0].rows);
    cols = std::move(M.staged[0].cols);
    vals = std::move(M.staged[0].vals);
    nu = M.staged[0].n;
  } else if (ntot > 0) {
    DVec<i64> r0((size_t)ntot), c0((size_t)ntot);
    DVec<double> v0((size_t)ntot);
    DVec<signed char> o0((size_t)ntot);
    i64 off = 0;
    for (auto &c : M.staged) {
      if (!c.n) continue;
      B200_CUDA(cudaMemcpyAsync(r0.p + off, c.rows.p, c.n * sizeof(i64), cudaMemcpyDeviceToDevice, stream()));
      B200_CUDA(cudaMemcpyAsync(c0.p + off, c.cols.p, c.n * sizeof(i64), cudaMemcpyDeviceToDevice, stream()));
      B200_CUDA(cudaMemcpyAsync(v0.p + off, c.vals.p, c.n * sizeof(double), cudaMemcpyDeviceToDevice, stream()));
      B200_CUDA(cudaMemsetAsync(o0.p + off, c.op, c.n, stream()));
      off += c.n;
    }
    // rows must lie in this rank's range before they are packed into sort keys
    bad.zero();
    k_check_range<<<div_up(ntot, 256), 256, 0, stream()>>>(ntot, r0.p, M.ilower, M.iupper, bad.p);
    B200_REQUIRE(fetch1(bad.p) == 0, "IJ assemble: a row index lies outside [ilower, iupper] of this rank");
    k_check_range<<<div_up(ntot, 256), 256, 0, stream()>>>(ntot, c0.p, 0, (i64)0x7fffffffffffffffLL, bad.p);
    B200_REQUIRE(fetch1(bad.p) == 0, "IJ assemble: negative column index");
    // stable sort by (row, col).  When (local row, column) fits one 64-bit key -- it always does for int32-sized ranks --
    // ONE radix sort over just the significant bits replaces the two full 64-bit passes of the LSD pair sort.
    DVec<int> perm0((size_t)ntot), perm1((size_t)ntot), perm2((size_t)ntot);
    DVec<i64> k1((size_t)ntot), k2((size_t)ntot), ktmp((size_t)ntot);
    k_iota64<<<div_up(ntot, 256), 256, 0, stream()>>>(ntot, perm0.p);
    i64 cmax = 0;
    {
      DVec<i64> dmax(1);
      size_t rb = 0;
      cub::DeviceReduce::Max(nullptr, rb, c0.p, dmax.p, (int)ntot, stream());
      DVec<char> rtmp(rb);
      cub::DeviceReduce::Max(rtmp.p, rb, c0.p, dmax.p, (int)ntot, stream());
      cmax = fetch1(dmax.p);
    }
    int cbits = 1, rbits = 1;
    while (cbits < 63 && (1LL << cbits) <= cmax) ++cbits;
    while (rbits < 63 && (1LL << rbits) <= (i64)nloc) ++rbits;
    size_t tb = 0;
    if (cmax >= 0 && rbits + cbits <= 63) {
      k_pack_keys<<<div_up(ntot, 256), 256, 0, stream()>>>(ntot, r0.p, c0.p, M.ilower, cbits, ktmp.p);
      cub::DeviceRadixSort::SortPairs(nullptr, tb, ktmp.p, k2.p, perm0.p, perm2.p, (int)ntot, 0, rbits + cbits, stream());
      DVec<char> tmp(tb);
      cub::DeviceRadixSort::SortPairs(tmp.p, tb, ktmp.p, k2.p, perm0.p, perm2.p, (int)ntot, 0, rbits + cbits, stream());
    } else {
      cub::DeviceRadixSort::SortPairs(nullptr, tb, c0.p, k1.p, perm0.p, perm1.p, (int)ntot, 0, 64, stream());
      DVec<char> tmp(tb);
      cub::DeviceRadixSort::SortPairs(tmp.p, tb, c0.p, k1.p, perm0.p, perm1.p, (int)ntot, 0, 64, stream());
      k_gather<i64><<<div_up(ntot, 256), 256, 0, stream()>>>(ntot, perm1.p, r0.p, ktmp.p);  // rows in col-sorted order
      cub::DeviceRadixSort::SortPairs(tmp.p, tb, ktmp.p, k2.p, perm1.p, perm2.p, (int)ntot, 0, 64, stream());
    }
    // gather everything by perm2
    DVec<i64> rs((size_t)ntot), cs((size_t)ntot);
    DVec<double> vs((size_t)ntot);
    DVec<signed char> os((size_t)ntot);
    k_gather<i64><<<div_up(ntot, 256), 256, 0, stream()>>>(ntot, perm2.p, r0.p, rs.p);
    k_gather<i64><<<div_up(ntot, 256), 256, 0, stream()>>>(ntot, perm2.p, c0.p, cs.p);
    k_gather<double><<<div_up(ntot, 256), 256, 0, stream()>>>(ntot, perm2.p, v0.p, vs.p);
    k_gather<signed char><<<div_up(ntot, 256), 256, 0, stream()>>>(ntot, perm2.p, o0.p, os.p);
    DVec<int> head((size_t)ntot + 1), slot((size_t)ntot + 1);
    head.zero();
    k_heads<<<div_up(ntot, 256), 256, 0, stream()>>>(ntot, rs.p, cs.p, head.p);
    exscan(head.p, slot.p, ntot + 1);
    nu = fetch1(slot.p + ntot);
    rows.alloc((size_t)nu);
    cols.alloc((size_t)nu);
    vals.alloc((size_t)nu);
    k_combine<<<div_up(ntot, 256), 256, 0, stream()>>>(ntot, head.p, slot.p, rs.p, cs.p, vs.p, os.p, rows.p, cols.p, vals.p);
    count_launch(12);
    B200_KERNEL_CHECK();
  }
  M.staged.clear();
  if (nu) {
    bad.zero();
    k_check_range<<<div_up(nu, 256), 256, 0, stream()>>>(nu, rows.p, M.ilower, M.iupper, bad.p);
    count_launch();
    B200_REQUIRE(fetch1(bad.p) == 0, "IJ assemble: a row index lies outside [ilower, iupper] of this rank");
  }
  // row pointers
  DVec<int> dcnt((size_t)nloc + 1), ocnt((size_t)nloc + 1);
  dcnt.zero();
  ocnt.zero();
  if (nu) k_row_counts<<<div_up(nu, 256), 256, 0, stream()>>>(nu, rows.p, cols.p, M.ilower, M.jlower, M.jupper, dcnt.p, ocnt.p);
  count_launch();
  A.diag.nr = nloc;
  A.diag.nc = A.n_cols_local;
  A.diag.ptr.alloc((size_t)nloc + 1);
  A.offd.nr = nloc;
  A.offd.ptr.alloc((size_t)nloc + 1);
  exscan(dcnt.p, A.diag.ptr.p, (i64)nloc + 1);
  exscan(ocnt.p, A.offd.ptr.p, (i64)nloc + 1);
  A.diag.nnz = fetch1(A.diag.ptr.p + nloc);
  A.offd.nnz = fetch1(A.offd.ptr.p + nloc);
  B200_REQUIRE(nranks > 1 || A.offd.nnz == 0, "IJ assemble: column index outside [jlower, jupper] on a single-rank run");
  A.diag.col.alloc((size_t)A.diag.nnz);
  A.diag.val.alloc((size_t)A.diag.nnz);
  A.offd.col.alloc((size_t)A.offd.nnz);
  A.offd.val.alloc((size_t)A.offd.nnz);
  // column map of the offd block: sorted unique global ids
  DVec<i64> colmap;
  A.col_map_offd.clear();
  if (A.offd.nnz > 0) {
    DVec<i64> oc((size_t)A.offd.nnz), ocs((size_t)A.offd.nnz), ocu((size_t)A.offd.nnz);
    DVec<int> counter(1), nsel(1);
    counter.zero();
    k_collect_offd<<<div_up(nu, 256), 256, 0, stream()>>>(nu, cols.p, M.jlower, M.jupper, oc.p, counter.p);
    size_t tb = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, tb, oc.p, ocs.p, (int)A.offd.nnz, 0, 64, stream());
    DVec<char> tmp(tb);
    cub::DeviceRadixSort::SortKeys(tmp.p, tb, oc.p, ocs.p, (int)A.offd.nnz, 0, 64, stream());
    size_t tb2 = 0;
    cub::DeviceSelect::Unique(nullptr, tb2, ocs.p, ocu.p, nsel.p, (int)A.offd.nnz, stream());
    DVec<char> tmp2(tb2);
    cub::DeviceSelect::Unique(tmp2.p, tb2, ocs.p, ocu.p, nsel.p, (int)A.offd.nnz, stream());
    int nh = fetch1(nsel.p);
    A.col_map_offd.resize(nh);
    B200_CUDA(cudaMemcpyAsync(A.col_map_offd.data(), ocu.p, nh * sizeof(i64), cudaMemcpyDeviceToHost, stream()));
    B200_CUDA(cudaStreamSynchronize(stream()));
    colmap.from_host(A.col_map_offd);
    count_launch(4);
  }
  A.n_halo = (int)A.col_map_offd.size();
  A.offd.nc = A.n_halo;
  // row starts of the sorted COO = diag.ptr + offd.ptr
  DVec<int> rowstart((size_t)nloc + 1);
  ij_add_int(nloc + 1, A.diag.ptr.p, A.offd.ptr.p, rowstart.p);
  if (nloc)
    k_split_fill<<<div_up(nloc, 256), 256, 0, stream()>>>(nloc, rowstart.p, cols.p, vals.p, M.ilower, M.jlower, M.jupper,
                                                          A.diag.ptr.p, A.diag.col.p, A.diag.val.p, A.offd.ptr.p, A.offd.col.p,
                                                          A.offd.val.p, colmap.p, A.n_halo);
  count_launch();
  // rows with offd entries
  if (A.offd.nnz > 0) {
    DVec<int> flag((size_t)nloc + 1), pos((size_t)nloc + 1);
    k_offd_rows<<<div_up((i64)nloc + 1, 256), 256, 0, stream()>>>(nloc, A.offd.ptr.p, flag.p);
    exscan(flag.p, pos.p, (i64)nloc + 1);
    int nr = fetch1(pos.p + nloc);
    A.offd_rows.alloc((size_t)nr);
    k_compact_rows<<<div_up(nloc, 256), 256, 0, stream()>>>(nloc, flag.p, pos.p, A.offd_rows.p);
    count_launch(2);
  }
  B200_KERNEL_CHECK();
  // global sizes / partition
  i64 g[2] = {(i64)nloc, (i64)A.n_cols_local};
  allreduce_sum_host_i64(g, 2);
  A.g_rows = g[0];
  A.g_cols = g[1];
  build_partition_and_halo(A);
  A.sell = DSell();
  A.finalize_spmv();
  M.assembled = true;
  B200_CUDA(cudaStreamSynchronize(stream()));
}

__global__ void k_add_int(int n, const int *a, const int *b, int *c) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) c[i] = a[i] + b[i];
}
void ij_add_int(int n, const int *a, const int *b, int *c) {
  if (n > 0) k_add_int<<<div_up(n, 256), 256, 0, stream()>>>(n, a, b, c);
  count_launch();
}

void ij_vector_init(IJVector &v) {
  ensure_init();
  int n = (int)(v.jupper - v.jlower + 1);
  v.par.alloc(n);
  v.par.start = v.jlower;
  i64 g = n;
  allreduce_sum_host_i64(&g, 1);
  v.par.g_n = g;
}

void ij_vector_set(IJVector &v, i64 n, const i64 *idx, const double *vals, int op) {
  ensure_init();
  if (n <= 0) return;
  B200_REQUIRE(v.par.p != nullptr, "IJVector: SetValues before Initialize");
  DVec<i64> di;
  DVec<double> dv;
  const i64 *pi = idx;
  const double *pv = vals;
  if (idx && !is_device_ptr(idx)) { di.from_host(idx, (size_t)n); pi = di.p; g_h2d_bytes += (unsigned long long)n * sizeof(i64); }
  if (!is_device_ptr(vals)) { dv.from_host(vals, (size_t)n); pv = dv.p; g_h2d_bytes += (unsigned long long)n * sizeof(double); }
  k_vec_set<<<div_up(n, 256), 256, 0, stream()>>>(n, pi, pv, v.jlower, v.jupper, v.par.p, op, v.jlower);
  count_launch();
  B200_CUDA(cudaStreamSynchronize(stream()));
}

void ij_vector_get(IJVector &v, i64 n, const i64 *idx, double *out) {
  ensure_init();
  if (n <= 0) return;
  DVec<i64> di;
  const i64 *pi = idx;
  if (idx && !is_device_ptr(idx)) { di.from_host(idx, (size_t)n); pi = di.p; }
  if (is_device_ptr(out)) {
    k_vec_get<<<div_up(n, 256), 256, 0, stream()>>>(n, pi, out, v.jlower, v.jupper, v.par.p);
    count_launch();
    B200_CUDA(cudaStreamSynchronize(stream()));
  } else if (!idx && n == v.par.n) {
    B200_CUDA(cudaMemcpyAsync(out, v.par.p, n * sizeof(double), cudaMemcpyDeviceToHost, stream()));
    B200_CUDA(cudaStreamSynchronize(stream()));
  } else {
    DVec<double> tmp((size_t)n);
    k_vec_get<<<div_up(n, 256), 256, 0, stream()>>>(n, pi, tmp.p, v.jlower, v.jupper, v.par.p);
    count_launch();
    B200_CUDA(cudaMemcpyAsync(out, tmp.p, n * sizeof(double), cudaMemcpyDeviceToHost, stream()));
    B200_CUDA(cudaStreamSynchronize(stream()));
  }
}

}  // namespace b200
