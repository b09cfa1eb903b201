// spmv.cu -- SpMV-class kernels (y = A x with fused epilogues) for sm_100a.
// Replaces hypre's ParCSR matvec / relaxation kernels reached from HYPRE_ParCSRGMRESSolve and HYPRE_BoomerAMGSolve
// (/root/reference/src/HypreSystem.cpp:723, :324; SURVEY.md §2.2 "ParCSR SpMV", A.7, A.8).
//
// Two layouts, both HBM-bound (12 B/nnz + 20 B/row algorithmic):
//   * SELL-32 (k_sell): thread per row, slices of 32 rows padded to the slice maximum; a thread reads two
//     consecutive entries of its row per iteration with one 128-bit (values) and one 64-bit (columns) load, the warp
//     reads 512 B + 256 B fully coalesced, and neighbouring threads gather neighbouring x entries (stencil matrices),
//     so x is served by L1/L2.  Used when padding is <= 20 %.
//   * CSR-vector (k_csr_vec<G>): G = 1..32 lanes per row chosen from the mean row length, warp-shuffle reduction.
// The epilogue (residual, l1-Jacobi update, prolongation add, Chebyshev Horner step) is fused so that each
// smoother sweep is exactly one pass over the matrix.
#include "common.h"
#include <cub/cub.cuh>

namespace b200 {

namespace {

template <int MODE>
__device__ __forceinline__ void epilogue(const SpmvArgs &a, int i, double t) {
  if (MODE == SPMV_AX) {
    a.y[i] = t;
  } else if (MODE == SPMV_RES) {
    double r = a.b[i] - t;
    if (a.d) r *= a.d[i];
    a.y[i] = r;
  } else if (MODE == SPMV_JAC) {
    a.y[i] = a.x[i] + a.a0 * (a.b[i] - t) / a.d[i];
  } else if (MODE == SPMV_ADD) {
    a.y[i] += t;
  } else {  // SPMV_CHEB
    double di = a.d[i];
    double r = a.a0 * a.b[i] + di * a.a1 * t;
    if (a.pt) r *= di;  // final Horner step: u += D^-1/2 v
    a.y[i] = (a.g ? a.g[i] : 0.0) + r;
  }
}

// U rows per lane group are in flight at once: a short row is a chain of three dependent loads (row pointer -> column /
// value -> x), so one row per group leaves the memory system idle two thirds of the time (P and R of the first levels:
// 2.5 - 3.2 TB/s with U = 1).  The U rows of a group are a block-stride apart, so every load phase stays coalesced.
template <int G, int MODE, bool XS, int U, bool IL>
__global__ void __launch_bounds__(256) k_csr_vec(CsrView A, SpmvArgs a) {
  constexpr int RPB = 256 / G;                       // rows per block and batch
  const int lane = threadIdx.x & (G - 1);
  const int r0 = blockIdx.x * (RPB * U) + (threadIdx.x / G);
  int row[U], kb[U], ke[U];
  bool live[U], masked[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    row[u] = r0 + u * RPB;
    live[u] = row[u] < A.nr;
    masked[u] = false;
    kb[u] = ke[u] = 0;
    if (live[u]) {
      if (MODE == SPMV_JAC && a.cf != nullptr) masked[u] = (a.cf[row[u]] != a.pt);
      if (!masked[u]) { kb[u] = __ldg(A.ptr + row[u]) + lane; ke[u] = __ldg(A.ptr + row[u] + 1); }
    }
  }
  int c[U];
  double v[U], sum[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    c[u] = 0; v[u] = 0.0; sum[u] = 0.0;
    if (kb[u] < ke[u]) { c[u] = __ldg(A.col + kb[u]); v[u] = __ldg(A.val + kb[u]); }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    if (kb[u] < ke[u]) {
      double xv = a.x[c[u]];
      if (XS) xv *= a.xs[c[u]];
      sum[u] = v[u] * xv;
    }
  }
  // rows longer than the group.  IL (matrices with long rows, e.g. the level-1 operator: 29 entries on 4 lanes = 7 more
  // rounds): the remaining rounds keep the U rows interleaved (load phase for all rows, then gather phase) like the first one;
  // measured -0.1 ms per GMRES iteration at 256^3.  Short rows (P, R of level 0: at most two more rounds) are faster row by row.
  if constexpr (IL) {
    bool more = false;
#pragma unroll
    for (int u = 0; u < U; ++u) { kb[u] += G; more |= kb[u] < ke[u]; }
    while (more) {
      more = false;
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (kb[u] < ke[u]) { c[u] = __ldg(A.col + kb[u]); v[u] = __ldg(A.val + kb[u]); }
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (kb[u] < ke[u]) {
          double xv = a.x[c[u]];
          if (XS) xv *= a.xs[c[u]];
          sum[u] += v[u] * xv;
          kb[u] += G;
          more |= kb[u] < ke[u];
        }
    }
  } else {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      for (int k = kb[u] + G; k < ke[u]; k += G) {
        int cc = __ldg(A.col + k);
        double vv = __ldg(A.val + k);
        double xv = a.x[cc];
        if (XS) xv *= a.xs[cc];
        sum[u] += vv * xv;
      }
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    if (G > 1) {
#pragma unroll
      for (int o = G >> 1; o > 0; o >>= 1) sum[u] += __shfl_down_sync(0xffffffffu, sum[u], o, G);
    }
    if (live[u] && lane == 0) {
      if (masked[u]) a.y[row[u]] = a.x[row[u]];
      else epilogue<MODE>(a, row[u], sum[u]);
    }
  }
}

template <int MODE, bool XS>
__global__ void __launch_bounds__(256) k_sell(int nr, const i64 *__restrict__ slice_off, const int *__restrict__ col,
                                              const double *__restrict__ val, SpmvArgs a) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  const int slice = row >> 5, lane = row & 31;
  if (row >= ((nr + 31) & ~31)) return;
  if (MODE == SPMV_JAC && a.cf != nullptr && row < nr) {
    if (a.cf[row] != a.pt) { a.y[row] = a.x[row]; return; }
  }
  const i64 off = slice_off[slice];
  const int pairs = (int)((slice_off[slice + 1] - off) >> 6);  // (entries per row) / 2
  const double2 *v2 = reinterpret_cast<const double2 *>(val + off) + lane;
  const int2 *c2 = reinterpret_cast<const int2 *>(col + off) + lane;
  double s0 = 0.0, s1 = 0.0;
  int k = 0;
  for (; k + 2 <= pairs; k += 2) {  // 2 x (128-bit + 64-bit) loads in flight per thread
    double2 va = __ldg(v2 + (k << 5));
    int2 ca = __ldg(c2 + (k << 5));
    double2 vb = __ldg(v2 + ((k + 1) << 5));
    int2 cb = __ldg(c2 + ((k + 1) << 5));
    double xa0 = a.x[ca.x], xa1 = a.x[ca.y], xb0 = a.x[cb.x], xb1 = a.x[cb.y];
    if (XS) { xa0 *= a.xs[ca.x]; xa1 *= a.xs[ca.y]; xb0 *= a.xs[cb.x]; xb1 *= a.xs[cb.y]; }
    s0 += va.x * xa0;
    s1 += va.y * xa1;
    s0 += vb.x * xb0;
    s1 += vb.y * xb1;
  }
  if (k < pairs) {
    double2 va = __ldg(v2 + (k << 5));
    int2 ca = __ldg(c2 + (k << 5));
    double xa0 = a.x[ca.x], xa1 = a.x[ca.y];
    if (XS) { xa0 *= a.xs[ca.x]; xa1 *= a.xs[ca.y]; }
    s0 += va.x * xa0;
    s1 += va.y * xa1;
  }
  if (row < nr) epilogue<MODE>(a, row, s0 + s1);
}

// ---- SELL construction ---------------------------------------------------------------------------------------
__global__ void k_sell_widths(int nr, const int *__restrict__ ptr, i64 *__restrict__ slice_entries) {
  int row = blockIdx.x * blockDim.x + threadIdx.x;
  int len = (row < nr) ? (ptr[row + 1] - ptr[row]) : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) len = max(len, __shfl_xor_sync(0xffffffffu, len, o));
  len = (len + 1) & ~1;
  if ((threadIdx.x & 31) == 0 && (row >> 5) < ((nr + 31) >> 5)) slice_entries[row >> 5] = (i64)len * 32;
}
__global__ void k_sell_fill(int nr, CsrView A, const i64 *__restrict__ slice_off, int *__restrict__ scol, double *__restrict__ sval) {
  int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= ((nr + 31) & ~31)) return;
  int slice = row >> 5, lane = row & 31;
  i64 off = slice_off[slice];
  int w = (int)((slice_off[slice + 1] - off) >> 5);
  int b = 0, e = 0;
  if (row < nr) { b = A.ptr[row]; e = A.ptr[row + 1]; }
  int padcol = (e > b) ? A.col[e - 1] : 0;
  for (int k = 0; k < w; ++k) {
    i64 idx = off + (i64)(k >> 1) * 64 + lane * 2 + (k & 1);
    bool in = (b + k) < e;
    scol[idx] = in ? A.col[b + k] : padcol;
    sval[idx] = in ? A.val[b + k] : 0.0;
  }
}

template <int G, int MODE>
void launch_csr(const DCsr &A, const SpmvArgs &a) {
  constexpr int U = (G <= 8) ? 4 : (G == 16 ? 2 : 1);
  const int grid = div_up((i64)A.nr, (i64)(256 / G) * U);
  const bool il = U > 1 && A.nr > 0 && (double)A.nnz > 3.0 * G * (double)A.nr;   // more than two rounds beyond the first
  if (il) {
    if (a.xs) k_csr_vec<G, MODE, true, U, true><<<grid, 256, 0, stream()>>>(A.view(), a);
    else k_csr_vec<G, MODE, false, U, true><<<grid, 256, 0, stream()>>>(A.view(), a);
  } else {
    if (a.xs) k_csr_vec<G, MODE, true, U, false><<<grid, 256, 0, stream()>>>(A.view(), a);
    else k_csr_vec<G, MODE, false, U, false><<<grid, 256, 0, stream()>>>(A.view(), a);
  }
}
template <int MODE>
void dispatch_csr(const DCsr &A, int lanes, const SpmvArgs &a) {
  switch (lanes) {
    case 1: launch_csr<1, MODE>(A, a); break;
    case 2: launch_csr<2, MODE>(A, a); break;
    case 4: launch_csr<4, MODE>(A, a); break;
    case 8: launch_csr<8, MODE>(A, a); break;
    case 16: launch_csr<16, MODE>(A, a); break;
    default: launch_csr<32, MODE>(A, a); break;
  }
}
template <int MODE>
void launch_sell(const DSell &S, const SpmvArgs &a) {
  int grid = div_up(((i64)S.nr + 31) & ~31LL, 256);
  if (a.xs) k_sell<MODE, true><<<grid, 256, 0, stream()>>>(S.nr, S.slice_off.p, S.col.p, S.val.p, a);
  else k_sell<MODE, false><<<grid, 256, 0, stream()>>>(S.nr, S.slice_off.p, S.col.p, S.val.p, a);
}

}  // namespace

void build_sell(const DCsr &A, DSell &S) {
  S.nr = A.nr;
  S.nslices = (A.nr + 31) / 32;
  if (S.nslices == 0) return;
  DVec<i64> entries((size_t)S.nslices + 1);
  entries.zero();
  int rows_padded = S.nslices * 32;
  k_sell_widths<<<div_up(rows_padded, 256), 256, 0, stream()>>>(A.nr, A.ptr.p, entries.p);
  count_launch();
  S.slice_off.alloc((size_t)S.nslices + 1);
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, entries.p, S.slice_off.p, S.nslices + 1, stream());
  DVec<char> tmp(tb);
  cub::DeviceScan::ExclusiveSum(tmp.p, tb, entries.p, S.slice_off.p, S.nslices + 1, stream());
  i64 total = 0;
  B200_CUDA(cudaMemcpyAsync(&total, S.slice_off.p + S.nslices, sizeof(i64), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  S.padded = total;
  S.col.alloc((size_t)total);
  S.val.alloc((size_t)total);
  k_sell_fill<<<div_up(rows_padded, 256), 256, 0, stream()>>>(A.nr, A.view(), S.slice_off.p, S.col.p, S.val.p);
  count_launch(2);
  B200_KERNEL_CHECK();
}

// padded size a SELL-32 copy would have, without building it
static i64 sell_padded_size(const DCsr &A) {
  int ns = (A.nr + 31) / 32;
  if (ns == 0) return 0;
  DVec<i64> entries((size_t)ns);
  k_sell_widths<<<div_up(ns * 32, 256), 256, 0, stream()>>>(A.nr, A.ptr.p, entries.p);
  count_launch();
  DVec<i64> total(1);
  size_t tb = 0;
  cub::DeviceReduce::Sum(nullptr, tb, entries.p, total.p, ns, stream());
  DVec<char> tmp(tb);
  cub::DeviceReduce::Sum(tmp.p, tb, entries.p, total.p, ns, stream());
  return total.to_host()[0];
}

SpmvPlan make_plan(const DCsr &A, const DSell *S) {
  SpmvPlan p;
  double mean = A.nr > 0 ? (double)A.nnz / A.nr : 0.0;
  int lanes = 1;
  const int per_lane = mean <= 12.0 ? 4 : 8;  // entries per lane: with 4 rows in flight per group, few lanes per row win (measured, r01zc/zd)
  while (lanes < 32 && lanes * per_lane <= mean) lanes *= 2;
  if (mean <= 3.0) lanes = 1;
  p.lanes = lanes;
  const char *force = getenv("B200_SPMV_KERNEL");  // "csr" | "sell" : benchmarking override
  p.use_sell = S && S->valid() && (double)S->padded <= 1.2 * (double)A.nnz;
  if (force && std::string(force) == "csr") p.use_sell = false;
  if (force && std::string(force) == "sell" && S && S->valid()) p.use_sell = true;
  const char *fl = getenv("B200_SPMV_LANES");
  if (fl && atoi(fl) > 0) p.lanes = atoi(fl);
  return p;
}

// decide whether a SELL copy is worth building (padding <= 20 %) and build it
void maybe_build_sell(const DCsr &A, DSell &S) {
  if (A.nr < 1024 || A.nnz == 0) return;
  const char *force = getenv("B200_SPMV_KERNEL");
  if (force && std::string(force) == "csr") return;
  i64 padded = sell_padded_size(A);
  bool ok = (double)padded <= 1.2 * (double)A.nnz;
  if (force && std::string(force) == "sell") ok = true;
  if (ok) build_sell(A, S);
}

// algorithmic bytes of one SpMV-class pass: 12 B/nnz + row pointer + y write + x read once, plus the
// epilogue's extra vectors; the same figure for both layouts (SELL padding is not counted as useful traffic)
static double spmv_bytes(const DCsr &A, SpmvMode mode, const SpmvArgs &a) {
  double nr = A.nr, nc = A.nc, b = 12.0 * (double)A.nnz + 12.0 * nr + 8.0 * nc;
  if (a.xs) b += 8.0 * nc;
  switch (mode) {
    case SPMV_AX: break;
    case SPMV_RES: b += 8.0 * nr + (a.d ? 8.0 * nr : 0.0); break;
    case SPMV_JAC: b += 16.0 * nr + (a.cf ? 4.0 * nr : 0.0); break;
    case SPMV_ADD: b += 8.0 * nr; break;
    case SPMV_CHEB: b += 16.0 * nr + (a.g ? 8.0 * nr : 0.0); break;
  }
  return b;
}

void spmv(const DCsr &A, const DSell *S, const SpmvPlan &plan, SpmvMode mode, const SpmvArgs &a) {
  if (A.nr == 0) return;
  static const char *const cls_sell[5] = {"spmv_sell32_ax", "spmv_sell32_res", "spmv_sell32_jac", "spmv_sell32_add", "spmv_sell32_cheb"};
  static const char *const cls_csr[5] = {"spmv_csrvec_ax", "spmv_csrvec_res", "spmv_csrvec_jac", "spmv_csrvec_add", "spmv_csrvec_cheb"};
  const bool sell = plan.use_sell && S && S->valid();
  ProfScope prof(sell ? cls_sell[mode] : cls_csr[mode], spmv_bytes(A, mode, a));
  if (sell) {
    switch (mode) {
      case SPMV_AX: launch_sell<SPMV_AX>(*S, a); break;
      case SPMV_RES: launch_sell<SPMV_RES>(*S, a); break;
      case SPMV_JAC: launch_sell<SPMV_JAC>(*S, a); break;
      case SPMV_ADD: launch_sell<SPMV_ADD>(*S, a); break;
      case SPMV_CHEB: launch_sell<SPMV_CHEB>(*S, a); break;
    }
  } else {
    switch (mode) {
      case SPMV_AX: dispatch_csr<SPMV_AX>(A, plan.lanes, a); break;
      case SPMV_RES: dispatch_csr<SPMV_RES>(A, plan.lanes, a); break;
      case SPMV_JAC: dispatch_csr<SPMV_JAC>(A, plan.lanes, a); break;
      case SPMV_ADD: dispatch_csr<SPMV_ADD>(A, plan.lanes, a); break;
      case SPMV_CHEB: dispatch_csr<SPMV_CHEB>(A, plan.lanes, a); break;
    }
  }
  count_launch();
  B200_KERNEL_CHECK();
}

}  // namespace b200
