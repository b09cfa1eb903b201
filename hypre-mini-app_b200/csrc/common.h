// common.h -- shared declarations of libb200hypre (B200 / sm_100a native GMRES + BoomerAMG path).
// Everything here is internal; the public surface is include/HYPRE*.h + include/b200hypre.h.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace b200 {

typedef long long i64;

[[noreturn]] void fail(const char *file, int line, const std::string &msg);

#define B200_CUDA(call)                                                               \
  do {                                                                                \
    cudaError_t e__ = (call);                                                         \
    if (e__ != cudaSuccess) ::b200::fail(__FILE__, __LINE__, cudaGetErrorString(e__)); \
  } while (0)
#define B200_REQUIRE(cond, msg)                                   \
  do {                                                            \
    if (!(cond)) ::b200::fail(__FILE__, __LINE__, std::string(msg)); \
  } while (0)
#define B200_KERNEL_CHECK() B200_CUDA(cudaGetLastError())

// ---- process-global context ------------------------------------------------------------------------------
struct Context {
  bool initialised = false;
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;       // every kernel of the library is launched here
  cudaStream_t comm_stream = nullptr;  // halo exchange
  int rank = 0, nranks = 1;
  void *nccl = nullptr;                // ncclComm_t when nranks > 1
  unsigned long long launches = 0;     // number of kernels launched by the library (gpu_launches in bench.py)
  unsigned long long rng_seed = 1234, rng_offset = 0;
};
Context &ctx();
void ensure_init();
inline cudaStream_t stream() { return ctx().stream; }
inline void count_launch(int n = 1) { ctx().launches += n; }

// ---- kernel-class profiler: CUDA-event pairs on stream() around the large kernels, summed per class name.  Enabled by
// b200hypre_profile_enable (bench.py's roofline leg); a scope is recorded only when its algorithmic byte count is at
// least the configured minimum and the stream is not being captured into a graph.
struct ProfScope {
  int slot = -1;
  ProfScope(const char *cls, double algorithmic_bytes);
  ~ProfScope();
};
void prof_enable(bool on, double min_bytes);
int prof_summary(std::vector<std::string> &names, std::vector<double> &ms, std::vector<double> &calls, std::vector<double> &bytes);

// ---- stream-ordered device buffer --------------------------------------------------------------------------
void *dev_alloc(size_t bytes);
void dev_free(void *p);

template <class T>
struct DVec {
  T *p = nullptr;
  size_t n = 0;
  DVec() {}
  explicit DVec(size_t n_) { alloc(n_); }
  DVec(const DVec &) = delete;
  DVec &operator=(const DVec &) = delete;
  DVec(DVec &&o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
  DVec &operator=(DVec &&o) noexcept {
    if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; }
    return *this;
  }
  ~DVec() { release(); }
  void alloc(size_t n_) {
    release();
    n = n_;
    p = n ? (T *)dev_alloc(n * sizeof(T)) : nullptr;
  }
  void release() {
    if (p) dev_free(p);
    p = nullptr;
    n = 0;
  }
  void zero() { if (n) B200_CUDA(cudaMemsetAsync(p, 0, n * sizeof(T), stream())); }
  void from_host(const T *h, size_t cnt) {
    if (cnt != n) alloc(cnt);
    if (cnt) B200_CUDA(cudaMemcpyAsync(p, h, cnt * sizeof(T), cudaMemcpyHostToDevice, stream()));
  }
  void from_host(const std::vector<T> &h) { from_host(h.data(), h.size()); }
  std::vector<T> to_host() const {
    std::vector<T> h(n);
    if (n) {
      B200_CUDA(cudaMemcpyAsync(h.data(), p, n * sizeof(T), cudaMemcpyDeviceToHost, stream()));
      B200_CUDA(cudaStreamSynchronize(stream()));
    }
    return h;
  }
  void swap(DVec &o) { std::swap(p, o.p); std::swap(n, o.n); }
};

inline int div_up(i64 a, i64 b) { return (int)((a + b - 1) / b); }

// ---- matrices -----------------------------------------------------------------------------------------------
struct CsrView {  // raw view usable in kernels (and in the host emulation of the setup row functions)
  int nr, nc;
  const int *ptr;
  const int *col;
  const double *val;
};

struct DCsr {  // device CSR block, int32 local indices, fp64 values
  int nr = 0, nc = 0;
  i64 nnz = 0;
  DVec<int> ptr, col;
  DVec<double> val;
  CsrView view() const { return CsrView{nr, nc, ptr.p, col.p, val.p}; }
};

// SELL-32 copy of a CSR block used by the SpMV-class kernels when padding is small.  Slice s covers rows
// [32 s, 32 s + 32); entry k of row r lives at slice_off[s] + (k/2)*64 + lane*2 + (k&1), so one thread reads two
// consecutive entries of its row with a single 128-bit (values) / 64-bit (columns) load.
struct DSell {
  int nr = 0, nslices = 0;
  i64 padded = 0;
  DVec<i64> slice_off;   // nslices + 1, in entries
  DVec<int> col;
  DVec<double> val;
  bool valid() const { return nslices > 0; }
};

// ---- SpMV-class operator ---------------------------------------------------------------------------------------
enum SpmvMode {
  SPMV_AX = 0,    // y = t                         t = sum_j a_ij x_j (xs: x_j scaled by xs_j)
  SPMV_RES = 1,   // y = rs_i (b_i - t)            rs optional
  SPMV_JAC = 2,   // y = x_i + a0 (b_i - t) / d_i  [rows with cf != pt copy x_i when cf given]
  SPMV_ADD = 3,   // y = y_i + t
  SPMV_CHEB = 4,  // y = g_i + p_i (a0 b_i + d_i a1 t), p_i = d_i if pt else 1  (g optional; Chebyshev Horner step)
};
struct SpmvArgs {
  const double *x = nullptr;   // gathered vector
  const double *xs = nullptr;  // optional gather scale
  double *y = nullptr;
  const double *b = nullptr;
  const double *d = nullptr;   // row diagonal / scale
  const double *g = nullptr;
  const int *cf = nullptr;
  int pt = 0;
  double a0 = 1.0, a1 = 1.0;
};

struct SpmvPlan {  // per-matrix kernel selection
  int lanes = 8;   // CSR-vector lanes per row
  bool use_sell = false;
};

void build_sell(const DCsr &A, DSell &S);
SpmvPlan make_plan(const DCsr &A, const DSell *S);
void spmv(const DCsr &A, const DSell *S, const SpmvPlan &plan, SpmvMode mode, const SpmvArgs &args);

// ---- vector kernels (vecops.cu) -----------------------------------------------------------------------------------
void vec_fill(double *x, double v, i64 n);
void vec_copy(double *dst, const double *src, i64 n);
void vec_scale(double *x, double a, i64 n);
void vec_axpy(double a, const double *x, double *y, i64 n);           // y += a x
void vec_axpby(double a, const double *x, double b, double *y, i64 n);  // y = a x + b y
void vec_pointwise_div(double *out, const double *f, const double *d, double w, const int *cf, int pt, i64 n);  // out = w f/d (masked rows: 0)
void vec_pointwise_mul(double *out, const double *a, const double *b, double c, i64 n);                       // out = c a.*b
// deterministic local dot products; results are left on the device (k doubles at out)
struct DotWork {
  DVec<double> partial;
  int blocks = 0;
};
void multi_dot(DotWork &w, const double *V, i64 ld, int k, const double *x, i64 n, double *out);  // out[j] = <V_j, x>, j<k
void multi_axpy_norm(DotWork &w, const double *V, i64 ld, int k, const double *h, double *x, i64 n, double *norm2_out);  // x -= sum h_j V_j ; norm2 = <x,x>
// x += a p ; r -= a s ; out[0] = <r,r>, out[1] = <t,r> of the updated r (t null: only out[0]; out null: no reduction)
void fused_update2(DotWork &w, double a, const double *p, const double *s, double *x, double *r, const double *t, i64 n, double *out);
void lin_comb(double *out, const double *V, i64 ld, int k, const double *coef_dev, i64 n);  // out = sum coef_j V_j
double dot_host(DotWork &w, const double *a, const double *b, i64 n);  // local + allreduce, returned on host

// ---- communication (comm.cpp); single rank: no-ops ---------------------------------------------------------------
void allreduce_sum_dev(double *buf, int count);     // in place on device buffer, ordered on stream()
void allreduce_sum_host(double *buf, int count);    // host values
void allreduce_sum_host_i64(i64 *buf, int count);
void allreduce_max_host_i64(i64 *buf, int count);
void barrier();

}  // namespace b200
