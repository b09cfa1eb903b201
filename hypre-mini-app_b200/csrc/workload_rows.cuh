// workload_rows.cuh -- row functions of the two synthetic nalu-wind-like systems of BASELINE.json configs[3] / configs[4]
// (SURVEY.md section 8d inputs 4 and 5), so that they can be generated ON THE DEVICE at their stated sizes (50 M / 100 M rows;
// the host builders of tests/workloads.py need minutes of numpy time and tens of GB for a tenth of that).
//   poisson27_hex : trilinear-hexahedron Galerkin Laplacian (2x2x2 Gauss) on the n^3 interior nodes of an (n+2)^3 grid with
//                   homogeneous Dirichlet boundary, interior nodes displaced by jitter * h * U(-1,1); 27-point, SPD, divided by h
//   momentum3     : n^3 nodes x 3 DOF, node-blocked, 7-point first-order upwind advection-diffusion at cell Peclet 2 with
//                   velocity (1, 0.5, 0.25) + 0.1 U(-1,1), a time term on the diagonal and a weak inter-component coupling
// Every row is a closed-form function of (n, seed, row id): the random numbers come from a counter-based generator
// (splitmix64 of seed and index), identical here and in tests/workloads.py, whose numpy builders stay the small-size
// oracle input.  These systems reach the path through the IJ interface like the reference's file-loaded systems
// (/root/reference/src/HypreSystem.cpp:897-1015).  __host__ __device__ so that tests/host/workload_rows.cu can compare them
// with the numpy builders in the GPU-less container; the product only runs them on the device (generator.cu).
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define B200_WL_HD __host__ __device__ __forceinline__
#else
#define B200_WL_HD inline
#endif

namespace b200 {
namespace wl {

// uniform in [-1, 1): splitmix64 finaliser of (seed, idx), top 53 bits
B200_WL_HD double uniform_pm1(unsigned long long seed, unsigned long long idx) {
  unsigned long long z = (seed + 1ULL) * 0x9E3779B97F4A7C15ULL + idx * 0xD1B54A32D192ED03ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  z = z ^ (z >> 31);
  return 2.0 * ((double)(z >> 11) * (1.0 / 9007199254740992.0)) - 1.0;
}

// ---- momentum3 ------------------------------------------------------------------------------------------------------------
// row = 3 * node + dof.  Writes the entries in ascending column order (at most 21), returns their number; *rhs = row sum.
B200_WL_HD int momentum3_row(int n, long long row, unsigned long long seed, long long *cols, double *vals, double *rhs) {
  const double peclet = 2.0, coupling = 0.05, time_term = 0.5;
  const double u0[3] = {1.0, 0.5, 0.25};
  const double M[3][3] = {{1.0, coupling, -coupling}, {-coupling, 1.0, coupling}, {coupling, -coupling, 1.0}};
  const double nu = sqrt(u0[0] * u0[0] + u0[1] * u0[1] + u0[2] * u0[2]) / peclet;   // h = 1
  const long long node = row / 3;
  const int p = (int)(row - 3 * node);
  const long long nn = (long long)n * n;
  const int c[3] = {(int)(node % n), (int)((node / n) % n), (int)(node / nn)};
  const long long stride[3] = {1, n, nn};
  double lo[3], hi[3], diag = time_term;
  for (int d = 0; d < 3; ++d) {
    const double u = u0[d] + 0.1 * uniform_pm1(seed, (unsigned long long)(node * 3 + d));
    lo[d] = nu + (u > 0.0 ? u : 0.0);      // coefficient magnitude on the (c - 1) neighbour: upwind for u > 0
    hi[d] = nu + (-u > 0.0 ? -u : 0.0);
    diag = diag + (lo[d] + hi[d]);
  }
  int cnt = 0;
  double sum = 0.0;
  // neighbour nodes in ascending id: -z, -y, -x, self, +x, +y, +z
  for (int t = 0; t < 7; ++t) {
    long long nb;
    double s;
    if (t < 3) { const int d = 2 - t; if (c[d] == 0) continue; nb = node - stride[d]; s = -lo[d]; }
    else if (t == 3) { nb = node; s = diag; }
    else { const int d = t - 4; if (c[d] == n - 1) continue; nb = node + stride[d]; s = -hi[d]; }
    for (int q = 0; q < 3; ++q) {
      const double v = s * M[p][q];
      cols[cnt] = 3 * nb + q;
      vals[cnt] = v;
      sum = sum + v;
      ++cnt;
    }
  }
  *rhs = sum;
  return cnt;
}

// ---- poisson27_hex --------------------------------------------------------------------------------------------------------
// grid node (I, J, K) of the (n+2)^3 grid, h = 1 / (n + 1); interior nodes 1 .. n are displaced
B200_WL_HD void hex_node_xyz(int n, int I, int J, int K, unsigned long long seed, double jitter, double h, double *x) {
  x[0] = I * h; x[1] = J * h; x[2] = K * h;
  if (I >= 1 && I <= n && J >= 1 && J <= n && K >= 1 && K <= n) {
    const unsigned long long id = (unsigned long long)((((long long)(K - 1) * n + (J - 1)) * n + (I - 1)) * 3);
    for (int d = 0; d < 3; ++d) x[d] += jitter * h * uniform_pm1(seed, id + d);
  }
}
// row of interior node `row` = ((k * n + j) * n + i), 0-based interior coordinates.  Entries in ascending column order
// (at most 27), returns their number; *rhs = row sum.
B200_WL_HD int poisson27_row(int n, long long row, unsigned long long seed, double jitter, long long *cols, double *vals, double *rhs) {
  const double h = 1.0 / (double)(n + 1);
  const long long nn = (long long)n * n;
  const int i = (int)(row % n), j = (int)((row / n) % n), k = (int)(row / nn);
  const int I = i + 1, J = j + 1, K = k + 1;
  // corner order of the element (xi, eta, zeta signs), as in tests/workloads.py
  const int cx[8] = {0, 1, 1, 0, 0, 1, 1, 0}, cy[8] = {0, 0, 1, 1, 0, 0, 1, 1}, cz[8] = {0, 0, 0, 0, 1, 1, 1, 1};
  const double g = 0.57735026918962576451;   // 1 / sqrt(3)
  double acc[27];
  for (int t = 0; t < 27; ++t) acc[t] = 0.0;
  for (int oz = 0; oz < 2; ++oz)
    for (int oy = 0; oy < 2; ++oy)
      for (int ox = 0; ox < 2; ++ox) {
        // element whose low corner is (I - ox, J - oy, K - oz); this node is its corner with offsets (ox, oy, oz)
        const int ei = I - ox, ej = J - oy, ek = K - oz;
        int la = 0;
        double X[8][3];
        for (int a = 0; a < 8; ++a) {
          hex_node_xyz(n, ei + cx[a], ej + cy[a], ek + cz[a], seed, jitter, h, X[a]);
          if (cx[a] == ox && cy[a] == oy && cz[a] == oz) la = a;
        }
        double krow[8];
        for (int b = 0; b < 8; ++b) krow[b] = 0.0;
        for (int q = 0; q < 8; ++q) {
          const double qx = (q & 4) ? g : -g, qy = (q & 2) ? g : -g, qz = (q & 1) ? g : -g;   // gx outermost, gz innermost
          double dN[8][3], Jm[3][3];
          for (int r = 0; r < 3; ++r)
            for (int s = 0; s < 3; ++s) Jm[r][s] = 0.0;
          for (int a = 0; a < 8; ++a) {
            const double sx = cx[a] ? 1.0 : -1.0, sy = cy[a] ? 1.0 : -1.0, sz = cz[a] ? 1.0 : -1.0;
            const double fx = 1.0 + sx * qx, fy = 1.0 + sy * qy, fz = 1.0 + sz * qz;
            dN[a][0] = sx * fy * fz / 8.0;
            dN[a][1] = sy * fx * fz / 8.0;
            dN[a][2] = sz * fx * fy / 8.0;
            for (int r = 0; r < 3; ++r)
              for (int s = 0; s < 3; ++s) Jm[r][s] += X[a][r] * dN[a][s];   // dx_r / dxi_s
          }
          const double c00 = Jm[1][1] * Jm[2][2] - Jm[1][2] * Jm[2][1], c01 = Jm[1][2] * Jm[2][0] - Jm[1][0] * Jm[2][2],
                       c02 = Jm[1][0] * Jm[2][1] - Jm[1][1] * Jm[2][0];
          const double det = Jm[0][0] * c00 + Jm[0][1] * c01 + Jm[0][2] * c02;
          const double id = 1.0 / det;
          double Ji[3][3];   // inverse of Jm
          Ji[0][0] = c00 * id;
          Ji[1][0] = c01 * id;
          Ji[2][0] = c02 * id;
          Ji[0][1] = (Jm[0][2] * Jm[2][1] - Jm[0][1] * Jm[2][2]) * id;
          Ji[1][1] = (Jm[0][0] * Jm[2][2] - Jm[0][2] * Jm[2][0]) * id;
          Ji[2][1] = (Jm[0][1] * Jm[2][0] - Jm[0][0] * Jm[2][1]) * id;
          Ji[0][2] = (Jm[0][1] * Jm[1][2] - Jm[0][2] * Jm[1][1]) * id;
          Ji[1][2] = (Jm[0][2] * Jm[1][0] - Jm[0][0] * Jm[1][2]) * id;
          Ji[2][2] = (Jm[0][0] * Jm[1][1] - Jm[0][1] * Jm[1][0]) * id;
          double G[8][3];   // dN_a / dx_r = sum_s dN[a][s] Ji[s][r]
          for (int a = 0; a < 8; ++a)
            for (int r = 0; r < 3; ++r) G[a][r] = dN[a][0] * Ji[0][r] + dN[a][1] * Ji[1][r] + dN[a][2] * Ji[2][r];
          for (int b = 0; b < 8; ++b) krow[b] += (G[la][0] * G[b][0] + G[la][1] * G[b][1] + G[la][2] * G[b][2]) * det;
        }
        for (int b = 0; b < 8; ++b) {
          const int di = cx[b] - ox, dj = cy[b] - oy, dk = cz[b] - oz;
          acc[(dk + 1) * 9 + (dj + 1) * 3 + (di + 1)] += krow[b];
        }
      }
  int cnt = 0;
  double sum = 0.0;
  for (int dk = -1; dk <= 1; ++dk)
    for (int dj = -1; dj <= 1; ++dj)
      for (int di = -1; di <= 1; ++di) {
        const int ii = i + di, jj = j + dj, kk = k + dk;
        if (ii < 0 || jj < 0 || kk < 0 || ii >= n || jj >= n || kk >= n) continue;   // Dirichlet nodes are eliminated
        const double v = acc[(dk + 1) * 9 + (dj + 1) * 3 + (di + 1)] / h;
        cols[cnt] = ((long long)kk * n + jj) * n + ii;
        vals[cnt] = v;
        sum = sum + v;
        ++cnt;
      }
  *rhs = sum;
  return cnt;
}

}  // namespace wl
}  // namespace b200
