// generator.cu -- 3-D Laplace weak-scaling generator on the device (CUDA; the reference's generator is HIP-only).
// Semantics of /root/reference/src/laplace_3d_weak_scaling.hpp (process grid :98-169, row/neighbour indexing :216-238,
// values :558,600, rhs :321) and of the glue in /root/reference/src/HypreSystem.cpp:1476-1607 (rank-contiguous global
// row ids :1516,1540): every rank owns an nx*ny*nz box of a (px*nx, py*ny, pz*nz) grid, rank -> box
// (rank % px, (rank / px) % py, rank / (px*py)); local rows are lexicographic (x fastest); global row =
// rank * nx*ny*nz + local; neighbours outside the global grid are dropped; diagonal = 26 (27-point, the reference)
// or 6 (7-point, BASELINE.json configs), off-diagonals -1, rhs = diag - (nnz_row - 1) so that x == 1.
// Off-rank neighbours carry the TRUE neighbour's global row id; apply_reference_halo_map (below) turns them into the reference
// glue's surrogate numbering (HypreSystem.cpp:1555) when the input asks for `halo_map: reference`.
// Output: COO triplets in row order, ascending global column inside a row (= stencil order z, y, x on one rank), plus the rhs.
#include "amg.h"
#include "workload_rows.cuh"
#include <cub/cub.cuh>

namespace b200 {

// process grid (laplace_3d_weak_scaling.hpp:80-169 restated): ascending prime factors f[0..m); m <= 3: the largest
// factor goes to x, the next to y, the smallest to z; m > 3: x starts from the largest factor and absorbs the smallest
// unused factors until it reaches cbrt(nprocs), y starts from the next largest and absorbs up to sqrt(nprocs / px),
// z takes what is left.  1 -> (1,1,1), 2 -> (2,1,1), 4 -> (2,2,1), 8 -> (2,2,2).
void process_grid(int nprocs, int *px, int *py, int *pz) {
  std::vector<int> f;
  for (int n = nprocs, d = 2; n > 1; ++d)
    for (; n % d == 0; n /= d) f.push_back(d);
  int g[3] = {1, 1, 1};
  const size_t m = f.size();
  if (m <= 3) {
    for (size_t k = 0; k < m; ++k) g[k] = f[m - 1 - k];
  } else {
    size_t lo = 0, hi = m;  // unused factors are f[lo, hi)
    g[0] = f[--hi];
    for (const double target = cbrt((double)nprocs); g[0] < target && lo < hi;) g[0] *= f[lo++];
    const double target_y = sqrt((double)(nprocs / g[0]));
    g[1] = f[--hi];
    while (g[1] < target_y && lo < hi) g[1] *= f[lo++];
    while (lo < hi) g[2] *= f[lo++];
  }
  B200_REQUIRE(g[0] * g[1] * g[2] == nprocs, "process_grid: factorisation does not cover nprocs");
  *px = g[0]; *py = g[1]; *pz = g[2];
}

namespace {

struct Box {
  int nx, ny, nz, px, py, pz, bx, by, bz, stencil;
  i64 nloc, row_start;
};

__device__ __forceinline__ bool in_stencil(int stencil, int dx, int dy, int dz) {
  return stencil == 27 || (abs(dx) + abs(dy) + abs(dz) <= 1);
}

__global__ void k_gen_count(Box b, int *cnt) {
  i64 l = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (l == b.nloc) cnt[l] = 0;
  if (l >= b.nloc) return;
  int lx = (int)(l % b.nx), ly = (int)((l / b.nx) % b.ny), lz = (int)(l / ((i64)b.nx * b.ny));
  int gx = b.bx * b.nx + lx, gy = b.by * b.ny + ly, gz = b.bz * b.nz + lz;
  int GX = b.nx * b.px, GY = b.ny * b.py, GZ = b.nz * b.pz;
  int c = 0;
  for (int dz = -1; dz <= 1; ++dz)
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx) {
        if (!in_stencil(b.stencil, dx, dy, dz)) continue;
        int x = gx + dx, y = gy + dy, z = gz + dz;
        if (x < 0 || y < 0 || z < 0 || x >= GX || y >= GY || z >= GZ) continue;
        ++c;
      }
  cnt[l] = c;
}

__global__ void k_gen_fill(Box b, const int *ptr, i64 *rows, i64 *cols, double *vals, double *rhs) {
  i64 l = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= b.nloc) return;
  int lx = (int)(l % b.nx), ly = (int)((l / b.nx) % b.ny), lz = (int)(l / ((i64)b.nx * b.ny));
  int gx = b.bx * b.nx + lx, gy = b.by * b.ny + ly, gz = b.bz * b.nz + lz;
  int GX = b.nx * b.px, GY = b.ny * b.py, GZ = b.nz * b.pz;
  const double dval = (b.stencil == 27) ? 26.0 : 6.0;
  i64 grow = b.row_start + l;
  int p = ptr[l], c = 0;
  i64 gc[27];
  for (int dz = -1; dz <= 1; ++dz)
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx) {
        if (!in_stencil(b.stencil, dx, dy, dz)) continue;
        int x = gx + dx, y = gy + dy, z = gz + dz;
        if (x < 0 || y < 0 || z < 0 || x >= GX || y >= GY || z >= GZ) continue;
        int rx = x / b.nx, ry = y / b.ny, rz = z / b.nz;
        i64 r = rx + (i64)b.px * (ry + (i64)b.py * rz);
        i64 gcol = r * b.nloc + (x - rx * b.nx) + (i64)b.nx * ((y - ry * b.ny) + (i64)b.ny * (z - rz * b.nz));
        // ascending global column inside the row (stencil order already is, except across rank boundaries), so that the
        // IJ assembly takes its pre-sorted path on every rank
        int q = c++;
        for (; q > 0 && gc[q - 1] > gcol; --q) gc[q] = gc[q - 1];
        gc[q] = gcol;
      }
  for (int q = 0; q < c; ++q) {
    rows[p + q] = grow;
    cols[p + q] = gc[q];
    vals[p + q] = (gc[q] == grow) ? dval : -1.0;
  }
  rhs[l] = dval - (double)(c - 1);
}

}  // namespace

void generate_laplace(int nx, int ny, int nz, int px, int py, int pz, int rank, int stencil, GenOut &out) {
  ensure_init();
  B200_REQUIRE(stencil == 7 || stencil == 27, "generator: stencil must be 7 or 27");
  B200_REQUIRE(nx > 0 && ny > 0 && nz > 0 && px > 0 && py > 0 && pz > 0 && rank >= 0 && rank < px * py * pz, "generator: bad box");
  Box b;
  b.nx = nx; b.ny = ny; b.nz = nz; b.px = px; b.py = py; b.pz = pz; b.stencil = stencil;
  b.bx = rank % px; b.by = (rank / px) % py; b.bz = rank / (px * py);
  b.nloc = (i64)nx * ny * nz;
  b.row_start = (i64)rank * b.nloc;
  B200_REQUIRE(b.nloc * stencil < (1LL << 31) - 8, "generator: more than 2^31 nonzeros per rank");
  DVec<int> cnt((size_t)b.nloc + 1), ptr((size_t)b.nloc + 1);
  int grid = div_up(b.nloc + 1, 256);
  k_gen_count<<<grid, 256, 0, stream()>>>(b, cnt.p);
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt.p, ptr.p, (int)(b.nloc + 1), stream());
  DVec<char> tmp(tb);
  cub::DeviceScan::ExclusiveSum(tmp.p, tb, cnt.p, ptr.p, (int)(b.nloc + 1), stream());
  int nnz = 0;
  B200_CUDA(cudaMemcpyAsync(&nnz, ptr.p + b.nloc, sizeof(int), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  out.nnz = nnz;
  out.n_local = b.nloc;
  out.row_start = b.row_start;
  out.g_rows = b.nloc * px * py * pz;
  out.rows.alloc((size_t)nnz);
  out.cols.alloc((size_t)nnz);
  out.vals.alloc((size_t)nnz);
  out.rhs.alloc((size_t)b.nloc);
  k_gen_fill<<<div_up(b.nloc, 256), 256, 0, stream()>>>(b, ptr.p, out.rows.p, out.cols.p, out.vals.p, out.rhs.p);
  count_launch(4);
  B200_KERNEL_CHECK();
  B200_CUDA(cudaStreamSynchronize(stream()));
}

// ---- `halo_map: reference`: the off-rank column numbering the reference glue really hands to hypre -----------------------------------
// /root/reference/src/HypreSystem.cpp:1553-1556 turns the generator's LOCAL halo index into a "global" column by adding iUpper + 1 on
// rank 0 and 0 on every other rank.  The generator numbers a rank's halo points 0 .. n_halo-1 by neighbour rank ascending and, inside a
// neighbour, by ascending grid id (src/laplace_3d_weak_scaling.hpp:495-531: slot = (dz+1)*9 + (dy+1)*3 + (dx+1); :966-1056: radix sort +
// run-length encode per neighbour, offsets accumulated in slot order) -- with rank-contiguous row ids that is the ascending order of the
// distinct off-rank columns of the rank.  So: column' = (position of the true column among the sorted distinct off-rank columns) +
// (rank == 0 ? i_upper + 1 : 0).  Every rank >= 1 couples to the first rows of rank 0, rank 0 to the first rows of rank 1: a
// nonsymmetric surrogate with the same row sums (x == 1 stays the solution).  In place on device COO columns; returns n_halo.
namespace {
__global__ void k_flag_offrank(i64 nnz, const i64 *__restrict__ cols, i64 lo, i64 hi, int *flag) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k <= nnz) flag[k] = (k < nnz && (cols[k] < lo || cols[k] > hi)) ? 1 : 0;
}
__global__ void k_gather_offrank(i64 nnz, const i64 *__restrict__ cols, const int *__restrict__ flag, const int *__restrict__ pos, i64 *out) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nnz && flag[k]) out[pos[k]] = cols[k];
}
__global__ void k_remap_offrank(i64 nnz, i64 *cols, const int *__restrict__ flag, const i64 *__restrict__ halo, int n_halo, i64 shift) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nnz || !flag[k]) return;
  const i64 c = cols[k];
  int lo = 0, hi = n_halo;   // lower bound
  while (lo < hi) {
    const int mid = lo + ((hi - lo) >> 1);
    if (halo[mid] < c) lo = mid + 1; else hi = mid;
  }
  cols[k] = (i64)lo + shift;
}
}  // namespace

i64 apply_reference_halo_map(i64 *cols, i64 nnz, i64 i_lower, i64 i_upper, int rank) {
  ensure_init();
  B200_REQUIRE(nnz >= 0 && nnz < (1LL << 31) - 8 && i_upper >= i_lower - 1, "halo_map: bad arguments");
  if (nnz == 0) return 0;
  DVec<int> flag((size_t)nnz + 1), pos((size_t)nnz + 1);
  const int grid = div_up(nnz + 1, 256);
  k_flag_offrank<<<grid, 256, 0, stream()>>>(nnz, cols, i_lower, i_upper, flag.p);
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, flag.p, pos.p, (int)(nnz + 1), stream());
  DVec<char> tmp(tb);
  cub::DeviceScan::ExclusiveSum(tmp.p, tb, flag.p, pos.p, (int)(nnz + 1), stream());
  int n_off = 0;
  B200_CUDA(cudaMemcpyAsync(&n_off, pos.p + nnz, sizeof(int), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  count_launch(3);
  if (n_off == 0) return 0;
  DVec<i64> off((size_t)n_off), sorted((size_t)n_off), uniq((size_t)n_off);
  DVec<int> n_uniq(1);
  k_gather_offrank<<<grid, 256, 0, stream()>>>(nnz, cols, flag.p, pos.p, off.p);
  cub::DeviceRadixSort::SortKeys(nullptr, tb, off.p, sorted.p, n_off, 0, 64, stream());
  DVec<char> tmp2(tb);
  cub::DeviceRadixSort::SortKeys(tmp2.p, tb, off.p, sorted.p, n_off, 0, 64, stream());
  cub::DeviceSelect::Unique(nullptr, tb, sorted.p, uniq.p, n_uniq.p, n_off, stream());
  DVec<char> tmp3(tb);
  cub::DeviceSelect::Unique(tmp3.p, tb, sorted.p, uniq.p, n_uniq.p, n_off, stream());
  int n_halo = 0;
  B200_CUDA(cudaMemcpyAsync(&n_halo, n_uniq.p, sizeof(int), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  k_remap_offrank<<<grid, 256, 0, stream()>>>(nnz, cols, flag.p, uniq.p, n_halo, rank == 0 ? i_upper + 1 : 0);
  count_launch(4);
  B200_KERNEL_CHECK();
  B200_CUDA(cudaStreamSynchronize(stream()));
  return n_halo;
}

// ---- synthetic nalu-wind-like systems (BASELINE.json configs[3] / [4]): rows [row_lo, row_hi] of the global system, built by
// the row functions of workload_rows.cuh.  COO triplets in row order with ascending columns (the pre-sorted path of the IJ
// assembly) and rhs = A 1.  kind 0: poisson27_hex(n), n^3 rows; kind 1: momentum3(n), 3 n^3 rows.
namespace {
__global__ void k_wl_count(int kind, int n, i64 row_lo, i64 nrows, int *cnt) {
  const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (t == nrows) cnt[t] = 0;
  if (t >= nrows) return;
  const i64 row = row_lo + t, node = kind == 0 ? row : row / 3, nn = (i64)n * n;
  const int c[3] = {(int)(node % n), (int)((node / n) % n), (int)(node / nn)};
  if (kind == 0) {
    int m = 1;
    for (int d = 0; d < 3; ++d) m *= 1 + (c[d] > 0 ? 1 : 0) + (c[d] < n - 1 ? 1 : 0);
    cnt[t] = m;
  } else {
    int m = 1;
    for (int d = 0; d < 3; ++d) m += (c[d] > 0 ? 1 : 0) + (c[d] < n - 1 ? 1 : 0);
    cnt[t] = 3 * m;
  }
}
__global__ void __launch_bounds__(128) k_wl_fill(int kind, int n, i64 row_lo, i64 nrows, unsigned long long seed, const int *ptr, i64 *rows,
                                                 i64 *cols, double *vals, double *rhs) {
  const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nrows) return;
  const i64 row = row_lo + t;
  long long c[27];
  double v[27], b;
  const int m = kind == 0 ? wl::poisson27_row(n, row, seed, 0.2, c, v, &b) : wl::momentum3_row(n, row, seed, c, v, &b);
  const int p = ptr[t];
  for (int k = 0; k < m; ++k) { rows[p + k] = row; cols[p + k] = c[k]; vals[p + k] = v[k]; }
  rhs[t] = b;
}
}  // namespace

void generate_workload(int kind, int n, i64 row_lo, i64 row_hi, unsigned long long seed, GenOut &out) {
  ensure_init();
  B200_REQUIRE(kind == 0 || kind == 1, "generate_workload: kind must be 0 (poisson27_hex) or 1 (momentum3)");
  const i64 g_rows = (kind == 0 ? 1 : 3) * (i64)n * n * n;
  B200_REQUIRE(n > 1 && row_lo >= 0 && row_hi < g_rows && row_hi >= row_lo - 1, "generate_workload: bad row range");
  const i64 nrows = row_hi - row_lo + 1;
  B200_REQUIRE(nrows * 27 < (1LL << 31) - 8 || kind == 1, "generate_workload: more than 2^31 nonzeros per rank");
  B200_REQUIRE(nrows * 21 < (1LL << 31) - 8 || kind == 0, "generate_workload: more than 2^31 nonzeros per rank");
  DVec<int> cnt((size_t)nrows + 1), ptr((size_t)nrows + 1);
  k_wl_count<<<div_up(nrows + 1, 256), 256, 0, stream()>>>(kind, n, row_lo, nrows, cnt.p);
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt.p, ptr.p, (int)(nrows + 1), stream());
  DVec<char> tmp(tb);
  cub::DeviceScan::ExclusiveSum(tmp.p, tb, cnt.p, ptr.p, (int)(nrows + 1), stream());
  int nnz = 0;
  B200_CUDA(cudaMemcpyAsync(&nnz, ptr.p + nrows, sizeof(int), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  out.nnz = nnz;
  out.n_local = nrows;
  out.row_start = row_lo;
  out.g_rows = g_rows;
  out.rows.alloc((size_t)std::max(nnz, 1));
  out.cols.alloc((size_t)std::max(nnz, 1));
  out.vals.alloc((size_t)std::max(nnz, 1));
  out.rhs.alloc((size_t)std::max<i64>(nrows, 1));
  if (nrows > 0) k_wl_fill<<<div_up(nrows, 128), 128, 0, stream()>>>(kind, n, row_lo, nrows, seed, ptr.p, out.rows.p, out.cols.p, out.vals.p, out.rhs.p);
  count_launch(4);
  B200_KERNEL_CHECK();
  B200_CUDA(cudaStreamSynchronize(stream()));
}

}  // namespace b200
