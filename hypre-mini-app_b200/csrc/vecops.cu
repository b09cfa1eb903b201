// vecops.cu -- vector kernels of the Krylov / V-cycle path: fill/copy/axpy, deterministic dot products and the
// fused Arnoldi kernels (multi-dot in one pass over the basis, fused update + norm).
// Replaces hypre's ParVector ops reached from HYPRE_ParCSRGMRESSolve (/root/reference/src/HypreSystem.cpp:723;
// SURVEY.md §2.2 "ParVector ops", A.1).  All kernels are HBM-bound streaming kernels: 128-bit loads
// (double2), grid = multiple of the SM count, warp-shuffle + shared-memory block reduction, per-block partials
// summed in a fixed order by the last block (deterministic, no floating-point atomics).
#include "common.h"
#include "abi_util.h"   // error-flag plumbing of the one C-ABI entry point defined here (b200hypre_multi_dot_noalloc)

namespace b200 {

namespace {

constexpr int VB = 256;  // threads per block for streaming kernels

inline int stream_grid(i64 n, int per_thread = 4) {
  i64 want = (n + (i64)VB * per_thread - 1) / ((i64)VB * per_thread);
  i64 cap = (i64)ctx().sm_count * 8;
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

__global__ void k_fill(double *x, double v, i64 n) {
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x, st = (i64)gridDim.x * blockDim.x;
  for (; i < n; i += st) x[i] = v;
}
__global__ void k_scale(double *x, double a, i64 n) {
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x, st = (i64)gridDim.x * blockDim.x;
  for (; i < n; i += st) x[i] *= a;
}
__global__ void k_axpy(double a, const double *__restrict__ x, double *__restrict__ y, i64 n) {
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x, st = (i64)gridDim.x * blockDim.x;
  for (; i < n; i += st) y[i] += a * x[i];
}
__global__ void k_axpby(double a, const double *__restrict__ x, double b, double *__restrict__ y, i64 n) {
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x, st = (i64)gridDim.x * blockDim.x;
  for (; i < n; i += st) y[i] = a * x[i] + b * y[i];
}
__global__ void k_pdiv(double *__restrict__ out, const double *__restrict__ f, const double *__restrict__ d, double w,
                       const int *__restrict__ cf, int pt, i64 n) {
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x, st = (i64)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    bool on = (cf == nullptr) || (cf[i] == pt);
    out[i] = on ? w * f[i] / d[i] : 0.0;
  }
}
__global__ void k_pmul(double *__restrict__ out, const double *__restrict__ a, const double *__restrict__ b, double c, i64 n) {
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x, st = (i64)gridDim.x * blockDim.x;
  for (; i < n; i += st) out[i] = c * a[i] * b[i];
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}

// Block-reduce K accumulators and store them to partial[block][K]; the last block to finish (ticket counter)
// sums the partials in block order and writes out[0..K).  `ticket` is reset by that block.
template <int K>
__device__ __forceinline__ void block_finish(double (&acc)[K], double *partial, double *out, unsigned int *ticket, int kvalid) {
  __shared__ double sm[K][VB / 32];
  __shared__ bool is_last;
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < K; ++j) {
    double v = warp_sum(acc[j]);
    if (lane == 0) sm[j][wid] = v;
  }
  __syncthreads();
  if (threadIdx.x < K) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < VB / 32; ++w) s += sm[threadIdx.x][w];
    partial[(size_t)blockIdx.x * K + threadIdx.x] = s;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    // all threads of the last block add the per-block partials (thread t takes blocks t, t + VB, ... in that order), then
    // the same fixed-order block reduction as above: deterministic for a given grid, and no serial chain of L2 loads
    __threadfence();
#pragma unroll
    for (int j = 0; j < K; ++j) {
      double s = 0.0;
      if (j < kvalid)
        for (unsigned int b = threadIdx.x; b < gridDim.x; b += VB) s += __ldcg(partial + (size_t)b * K + j);
      s = warp_sum(s);
      if (lane == 0) sm[j][wid] = s;
    }
    __syncthreads();
    if (threadIdx.x < kvalid) {
      double s = 0.0;
#pragma unroll
      for (int w = 0; w < VB / 32; ++w) s += sm[threadIdx.x][w];
      out[threadIdx.x] = s;
    }
    if (threadIdx.x == 0) *ticket = 0;
  }
}

// out[j] = <V_{j0+j}, x> for j < kvalid <= K : one pass over x per K basis vectors
template <int K>
__global__ void __launch_bounds__(VB) k_multi_dot(const double *__restrict__ V, i64 ld, int kvalid, const double *__restrict__ x, i64 n,
                                                  double *partial, double *out, unsigned int *ticket) {
  double acc[K];
#pragma unroll
  for (int j = 0; j < K; ++j) acc[j] = 0.0;
  i64 n2 = n >> 1;
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x, st = (i64)gridDim.x * blockDim.x;
  const double2 *x2 = reinterpret_cast<const double2 *>(x);
  for (; i < n2; i += st) {
    double2 xv = x2[i];
#pragma unroll
    for (int j = 0; j < K; ++j) {
      if (j < kvalid) {
        double2 vv = reinterpret_cast<const double2 *>(V + (i64)j * ld)[i];
        acc[j] += vv.x * xv.x + vv.y * xv.y;
      }
    }
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    double xv = x[n - 1];
#pragma unroll
    for (int j = 0; j < K; ++j)
      if (j < kvalid) acc[j] += V[(i64)j * ld + n - 1] * xv;
  }
  block_finish<K>(acc, partial, out, ticket, kvalid);
}

// x -= sum_j h_j V_j (j < k, any k), and norm2 = <x,x> of the result
__global__ void __launch_bounds__(VB) k_multi_axpy_norm(const double *__restrict__ V, i64 ld, int k, const double *__restrict__ h,
                                                        double *__restrict__ x, i64 n, double *partial, double *out, unsigned int *ticket) {
  __shared__ double hs[128];
  for (int j = threadIdx.x; j < k; j += blockDim.x) hs[j] = h[j];
  __syncthreads();
  double acc[1] = {0.0};
  i64 n2 = n >> 1;
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x, st = (i64)gridDim.x * blockDim.x;
  double2 *x2 = reinterpret_cast<double2 *>(x);
  for (; i < n2; i += st) {
    double2 xv = x2[i];
    int j = 0;
    for (; j + 4 <= k; j += 4) {  // 4 independent 128-bit loads in flight
      double2 v0 = reinterpret_cast<const double2 *>(V + (i64)(j + 0) * ld)[i];
      double2 v1 = reinterpret_cast<const double2 *>(V + (i64)(j + 1) * ld)[i];
      double2 v2 = reinterpret_cast<const double2 *>(V + (i64)(j + 2) * ld)[i];
      double2 v3 = reinterpret_cast<const double2 *>(V + (i64)(j + 3) * ld)[i];
      xv.x -= hs[j] * v0.x; xv.y -= hs[j] * v0.y;
      xv.x -= hs[j + 1] * v1.x; xv.y -= hs[j + 1] * v1.y;
      xv.x -= hs[j + 2] * v2.x; xv.y -= hs[j + 2] * v2.y;
      xv.x -= hs[j + 3] * v3.x; xv.y -= hs[j + 3] * v3.y;
    }
    for (; j < k; ++j) {
      double2 v0 = reinterpret_cast<const double2 *>(V + (i64)j * ld)[i];
      xv.x -= hs[j] * v0.x; xv.y -= hs[j] * v0.y;
    }
    x2[i] = xv;
    acc[0] += xv.x * xv.x + xv.y * xv.y;
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    double xv = x[n - 1];
    for (int j = 0; j < k; ++j) xv -= hs[j] * V[(i64)j * ld + n - 1];
    x[n - 1] = xv;
    acc[0] += xv * xv;
  }
  block_finish<1>(acc, partial, out, ticket, 1);
}

// Short-recurrence update of PCG / BiCGSTAB in one pass: x += a p ; r -= a s ; then, with the updated r,
// out[0] = <r,r> and out[1] = <t,r> (NDOT = 0: no reduction, 1: <r,r>, 2: both).
template <int NDOT>
__global__ void __launch_bounds__(VB) k_update2(double a, const double *__restrict__ p, const double *__restrict__ s, double *__restrict__ x,
                                                double *__restrict__ r, const double *__restrict__ t, i64 n, double *partial, double *out,
                                                unsigned int *ticket) {
  double acc[2] = {0.0, 0.0};
  i64 n2 = n >> 1;
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x, st = (i64)gridDim.x * blockDim.x;
  for (; i < n2; i += st) {
    double2 pv = reinterpret_cast<const double2 *>(p)[i], sv = reinterpret_cast<const double2 *>(s)[i];
    double2 xv = reinterpret_cast<double2 *>(x)[i], rv = reinterpret_cast<double2 *>(r)[i];
    xv.x += a * pv.x; xv.y += a * pv.y;
    rv.x -= a * sv.x; rv.y -= a * sv.y;
    reinterpret_cast<double2 *>(x)[i] = xv;
    reinterpret_cast<double2 *>(r)[i] = rv;
    if (NDOT >= 1) acc[0] += rv.x * rv.x + rv.y * rv.y;
    if (NDOT >= 2) {
      double2 tv = reinterpret_cast<const double2 *>(t)[i];
      acc[1] += tv.x * rv.x + tv.y * rv.y;
    }
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    double xv = x[n - 1] + a * p[n - 1], rv = r[n - 1] - a * s[n - 1];
    x[n - 1] = xv;
    r[n - 1] = rv;
    if (NDOT >= 1) acc[0] += rv * rv;
    if (NDOT >= 2) acc[1] += t[n - 1] * rv;
  }
  if (NDOT >= 1) block_finish<2>(acc, partial, out, ticket, NDOT);
}

// out = sum_j coef_j V_j
__global__ void __launch_bounds__(VB) k_lin_comb(double *__restrict__ out, const double *__restrict__ V, i64 ld, int k,
                                                 const double *__restrict__ coef, i64 n) {
  __shared__ double cs[128];
  for (int j = threadIdx.x; j < k; j += blockDim.x) cs[j] = coef[j];
  __syncthreads();
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x, st = (i64)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    double s = 0.0;
    for (int j = 0; j < k; ++j) s += cs[j] * V[(i64)j * ld + i];
    out[i] = s;
  }
}

unsigned int *ticket_ptr() {
  static unsigned int *t = nullptr;
  if (!t) {
    B200_CUDA(cudaMalloc(&t, sizeof(unsigned int)));
    B200_CUDA(cudaMemset(t, 0, sizeof(unsigned int)));
  }
  return t;
}

void ensure_work(DotWork &w, int blocks, int k) {
  size_t need = (size_t)blocks * k;
  if (w.partial.n < need) w.partial.alloc(need > 8192 ? need : 8192);
}

}  // namespace

void vec_fill(double *x, double v, i64 n) {
  if (n <= 0) return;
  if (v == 0.0) { B200_CUDA(cudaMemsetAsync(x, 0, n * sizeof(double), stream())); return; }
  k_fill<<<stream_grid(n), VB, 0, stream()>>>(x, v, n);
  count_launch();
}
void vec_copy(double *dst, const double *src, i64 n) {
  if (n > 0 && dst != src) B200_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(double), cudaMemcpyDeviceToDevice, stream()));
}
void vec_scale(double *x, double a, i64 n) {
  if (n <= 0) return;
  ProfScope prof("scale", 16.0 * (double)n);
  k_scale<<<stream_grid(n), VB, 0, stream()>>>(x, a, n);
  count_launch();
}
void vec_axpy(double a, const double *x, double *y, i64 n) {
  if (n <= 0) return;
  ProfScope prof("axpy", 24.0 * (double)n);
  k_axpy<<<stream_grid(n), VB, 0, stream()>>>(a, x, y, n);
  count_launch();
}
void vec_axpby(double a, const double *x, double b, double *y, i64 n) {
  if (n <= 0) return;
  k_axpby<<<stream_grid(n), VB, 0, stream()>>>(a, x, b, y, n);
  count_launch();
}
void vec_pointwise_div(double *out, const double *f, const double *d, double w, const int *cf, int pt, i64 n) {
  if (n <= 0) return;
  ProfScope prof("pointwise_div", 24.0 * (double)n);
  k_pdiv<<<stream_grid(n), VB, 0, stream()>>>(out, f, d, w, cf, pt, n);
  count_launch();
}
void vec_pointwise_mul(double *out, const double *a, const double *b, double c, i64 n) {
  if (n <= 0) return;
  k_pmul<<<stream_grid(n), VB, 0, stream()>>>(out, a, b, c, n);
  count_launch();
}

void multi_dot(DotWork &w, const double *V, i64 ld, int k, const double *x, i64 n, double *out) {
  B200_REQUIRE((ld & 1) == 0 && ((uintptr_t)V & 15) == 0 && ((uintptr_t)x & 15) == 0, "multi_dot: unaligned vectors");
  int grid = stream_grid(n > 0 ? n : 1, 8);
  ensure_work(w, grid, 8);
  unsigned int *t = ticket_ptr();
  ProfScope prof("multi_dot", 8.0 * (double)n * (k + (k + 7) / 8));
  for (int j0 = 0; j0 < k; j0 += 8) {
    int kv = (k - j0 < 8) ? (k - j0) : 8;
    const double *Vj = V + (i64)j0 * ld;
    if (kv <= 1) k_multi_dot<1><<<grid, VB, 0, stream()>>>(Vj, ld, kv, x, n, w.partial.p, out + j0, t);
    else if (kv <= 2) k_multi_dot<2><<<grid, VB, 0, stream()>>>(Vj, ld, kv, x, n, w.partial.p, out + j0, t);
    else if (kv <= 4) k_multi_dot<4><<<grid, VB, 0, stream()>>>(Vj, ld, kv, x, n, w.partial.p, out + j0, t);
    else k_multi_dot<8><<<grid, VB, 0, stream()>>>(Vj, ld, kv, x, n, w.partial.p, out + j0, t);
    count_launch();
  }
  B200_KERNEL_CHECK();
}

void multi_axpy_norm(DotWork &w, const double *V, i64 ld, int k, const double *h, double *x, i64 n, double *norm2_out) {
  B200_REQUIRE(k <= 128, "multi_axpy_norm: k > 128");
  B200_REQUIRE((ld & 1) == 0 && ((uintptr_t)V & 15) == 0 && ((uintptr_t)x & 15) == 0, "multi_axpy_norm: unaligned vectors");
  int grid = stream_grid(n > 0 ? n : 1, 8);
  ensure_work(w, grid, 8);
  ProfScope prof("multi_axpy_norm", 8.0 * (double)n * (k + 2));
  k_multi_axpy_norm<<<grid, VB, 0, stream()>>>(V, ld, k, h, x, n, w.partial.p, norm2_out, ticket_ptr());
  count_launch();
  B200_KERNEL_CHECK();
}

void fused_update2(DotWork &w, double a, const double *p, const double *s, double *x, double *r, const double *t, i64 n, double *out) {
  auto al = [](const void *q) { return ((uintptr_t)q & 15) == 0; };
  B200_REQUIRE(al(p) && al(s) && al(x) && al(r) && al(t), "fused_update2: unaligned vectors");
  int ndot = out ? (t ? 2 : 1) : 0;
  int grid = stream_grid(n > 0 ? n : 1, 8);
  ensure_work(w, grid, 8);
  ProfScope prof("krylov_update2", 8.0 * (double)n * (6 + (ndot == 2)));
  if (ndot == 0) k_update2<0><<<grid, VB, 0, stream()>>>(a, p, s, x, r, t, n, w.partial.p, out, ticket_ptr());
  else if (ndot == 1) k_update2<1><<<grid, VB, 0, stream()>>>(a, p, s, x, r, t, n, w.partial.p, out, ticket_ptr());
  else k_update2<2><<<grid, VB, 0, stream()>>>(a, p, s, x, r, t, n, w.partial.p, out, ticket_ptr());
  count_launch();
  B200_KERNEL_CHECK();
}

void lin_comb(double *out, const double *V, i64 ld, int k, const double *coef_dev, i64 n) {
  B200_REQUIRE(k <= 128, "lin_comb: k > 128");
  if (n <= 0) return;
  ProfScope prof("lin_comb", 8.0 * (double)n * (k + 1));
  k_lin_comb<<<stream_grid(n), VB, 0, stream()>>>(out, V, ld, k, coef_dev, n);
  count_launch();
  B200_KERNEL_CHECK();
}

double dot_host(DotWork &w, const double *a, const double *b, i64 n) {
  static double *d_out = nullptr;
  if (!d_out) B200_CUDA(cudaMalloc(&d_out, 16 * sizeof(double)));
  multi_dot(w, a, (n + 1) & ~1LL, 1, b, n, d_out);
  allreduce_sum_dev(d_out, 1);
  double h = 0.0;
  B200_CUDA(cudaMemcpyAsync(&h, d_out, sizeof(double), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  return h;
}

// ---- multi-dot that allocates nothing (b200hypre_multi_dot_noalloc, include/b200hypre.h) ------------------------------------------
// Same kernel and same grid as multi_dot above; the per-block partials and the ticket word live in a file-scope device array
// instead of a DotWork buffer, so that callers that cannot reach cudaMalloc (stream capture, serving loops) can use it.
// Every GPU that loads the library holds these NOALLOC_BYTES until the process exits.
namespace {
constexpr int NOALLOC_MAX_BLOCKS = 2048;                                            // launch grid the partials have room for
constexpr int NOALLOC_MAX_K = 64;                                                   // vectors per call (8 per pass over x)
constexpr size_t NOALLOC_PARTIAL_BYTES = (size_t)NOALLOC_MAX_BLOCKS * 8 * sizeof(double);   // partial[block][8]
constexpr size_t NOALLOC_BYTES = NOALLOC_PARTIAL_BYTES + 256;                       // + the ticket word in a line of its own
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];
// In host code g_noalloc_scratch is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress, once per
// device, with that device current.
constexpr int NOALLOC_MAX_DEVICES = 64;
unsigned char *g_noalloc_addr[NOALLOC_MAX_DEVICES] = {};

unsigned char *noalloc_scratch_on_current_device() {
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "multi_dot_noalloc: device ordinal beyond 63");
  if (!g_noalloc_addr[dev]) {
    void *p = nullptr;
    B200_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    B200_REQUIRE(p != nullptr, "multi_dot_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_addr[dev] = static_cast<unsigned char *>(p);
  }
  return g_noalloc_addr[dev];
}
}  // namespace

// out[j] = <V_j, x> over the n local entries, j < k; V_j = V + j ld.  Device pointers; ordered on stream(), not synchronised.
void multi_dot_noalloc(const double *V, i64 ld, int k, const double *x, i64 n, double *out) {
  B200_REQUIRE(ctx().initialised, "multi_dot_noalloc: library not initialised (HYPRE_Init creates the stream; this call creates nothing)");
  B200_REQUIRE(k >= 1 && k <= NOALLOC_MAX_K, "multi_dot_noalloc: k = " + std::to_string(k) + " vectors, the static scratch serves 1 .. " +
                                                 std::to_string(NOALLOC_MAX_K) + " per call");
  B200_REQUIRE(n >= 0 && (k == 1 || ld >= n), "multi_dot_noalloc: n < 0 or ld < n");
  B200_REQUIRE((k == 1 || (ld & 1) == 0) && ((uintptr_t)V & 15) == 0 && ((uintptr_t)x & 15) == 0,
               "multi_dot_noalloc: vectors must be 16-byte aligned and ld even");
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev == ctx().device, "multi_dot_noalloc: the current device is not the one the library was initialised on");
  const int grid = stream_grid(n > 0 ? n : 1, 8);
  B200_REQUIRE(grid <= NOALLOC_MAX_BLOCKS, "multi_dot_noalloc: launch grid of " + std::to_string(grid) + " blocks (n = " + std::to_string(n) +
                                               ") exceeds the " + std::to_string(NOALLOC_MAX_BLOCKS) + " blocks the static scratch has partials for");
  unsigned char *scr = noalloc_scratch_on_current_device();
  double *partial = reinterpret_cast<double *>(scr);
  unsigned int *ticket = reinterpret_cast<unsigned int *>(scr + NOALLOC_PARTIAL_BYTES);
  // the array keeps what the last call left: the ticket must start at zero (a call that died mid-kernel may have left a count);
  // the partials are written before they are read
  B200_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), stream()));
  for (int j0 = 0; j0 < k; j0 += 8) {
    const int kv = (k - j0 < 8) ? (k - j0) : 8;
    const double *Vj = V + (i64)j0 * ld;
    if (kv <= 1) k_multi_dot<1><<<grid, VB, 0, stream()>>>(Vj, ld, kv, x, n, partial, out + j0, ticket);
    else if (kv <= 2) k_multi_dot<2><<<grid, VB, 0, stream()>>>(Vj, ld, kv, x, n, partial, out + j0, ticket);
    else if (kv <= 4) k_multi_dot<4><<<grid, VB, 0, stream()>>>(Vj, ld, kv, x, n, partial, out + j0, ticket);
    else k_multi_dot<8><<<grid, VB, 0, stream()>>>(Vj, ld, kv, x, n, partial, out + j0, ticket);
    count_launch();
  }
  B200_KERNEL_CHECK();
}

}  // namespace b200

extern "C" HYPRE_Int b200hypre_multi_dot_noalloc(const HYPRE_Complex *V, HYPRE_BigInt ld, HYPRE_Int k, const HYPRE_Complex *x, HYPRE_BigInt n,
                                                 HYPRE_Complex *out) {
  API_ARG(V && x && out, "b200hypre_multi_dot_noalloc: null pointer");
  API_BEGIN
  b200::multi_dot_noalloc(V, ld, k, x, n, out);
  API_END
}
