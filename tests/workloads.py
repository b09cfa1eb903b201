"""Synthetic matrices of BASELINE.json configs[3] and configs[4] (SURVEY.md §8d inputs 4 and 5), scaled by `n`.

TEST / BENCH INPUT GENERATORS ONLY (numpy + scipy on the host); the product reads these systems through the IJ path
(HypreSystem.cpp:897-1015 semantics) or from MatrixMarket / HYPRE-IJ text written by `write_matrix_market` /
`write_hypre_ij` (formats: SURVEY.md Appendix C).

* poisson27_hex(n): nalu-wind-like pressure Poisson.  Trilinear-hexahedron Galerkin Laplacian (2x2x2 Gauss) on the
  n^3 interior nodes of an (n+2)^3 grid with homogeneous Dirichlet boundary; interior nodes displaced by
  0.2 h U(-1,1) (counter-based generator, seed 20260).  27-point stencil, nnz = (3n-2)^3, symmetric positive definite, positive and
  negative off-diagonals.  rhs = A 1.
* momentum3(n): nalu-wind-like momentum.  n^3 nodes x 3 DOF, node-blocked 3x3, 7-point first-order upwind
  advection-diffusion at cell Peclet 2 with velocity (1, 0.5, 0.25) + 0.1 U(-1,1) (counter-based generator, seed 20261), a time term
  on the diagonal, and a weak constant inter-component coupling.  Nonsymmetric, nnz = 9 (7 n^3 - 6 n^2).  rhs = A 1.
"""
import numpy as np
import scipy.sparse as sp


def uniform_pm1(seed, idx):
    """U[-1, 1) as a pure function of (seed, index): splitmix64 finaliser, top 53 bits.  The same generator as
    hypre-mini-app_b200/csrc/workload_rows.cuh (wl::uniform_pm1), so that the device generators of these systems and the numpy
    builders below produce the same matrices."""
    idx = np.asarray(idx, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = np.full(idx.shape, (int(seed) + 1) * 0x9E3779B97F4A7C15 % 2 ** 64, dtype=np.uint64) + idx * np.uint64(0xD1B54A32D192ED03)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return 2.0 * ((z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)) - 1.0


def _hex_element_matrices(X):
    """X: (ne, 8, 3) node coordinates in the usual (xi, eta, zeta) corner order -> (ne, 8, 8) stiffness matrices"""
    s = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]], dtype=np.float64)
    g = 1.0 / np.sqrt(3.0)
    K = np.zeros((X.shape[0], 8, 8))
    for gx in (-g, g):
        for gy in (-g, g):
            for gz in (-g, g):
                q = np.array([gx, gy, gz])
                f = 1.0 + s * q  # (8, 3)
                dN = np.empty((8, 3))
                dN[:, 0] = s[:, 0] * f[:, 1] * f[:, 2] / 8.0
                dN[:, 1] = s[:, 1] * f[:, 0] * f[:, 2] / 8.0
                dN[:, 2] = s[:, 2] * f[:, 0] * f[:, 1] / 8.0
                J = np.einsum("eai,aj->eij", X, dN)  # dx_i / dxi_j
                det = np.linalg.det(J)
                Ji = np.linalg.inv(J)
                G = np.einsum("aj,eji->eai", dN, Ji)  # dN_a / dx_i
                K += np.einsum("eai,ebi,e->eab", G, G, det)
    return K


def poisson27_hex(n, seed=20260, jitter=0.2):
    m = n + 2
    h = 1.0 / (m - 1)
    ax = np.arange(m) * h
    Z, Y, Xc = np.meshgrid(ax, ax, ax, indexing="ij")  # node id = (k * m + j) * m + i
    xyz = np.stack([Xc, Y, Z], axis=-1)
    disp = jitter * h * uniform_pm1(seed, np.arange(3 * n ** 3)).reshape(n, n, n, 3)   # [k][j][i][xyz]
    xyz[1:-1, 1:-1, 1:-1, :] += disp
    xyz = xyz.reshape(-1, 3)
    e = np.arange(m - 1)
    ek, ej, ei = np.meshgrid(e, e, e, indexing="ij")
    base = ((ek * m + ej) * m + ei).ravel()
    off = np.array([0, 1, m + 1, m, m * m, m * m + 1, m * m + m + 1, m * m + m])
    conn = base[:, None] + off[None, :]
    K = _hex_element_matrices(xyz[conn])
    rows = np.repeat(conn, 8, axis=1).ravel()
    cols = np.tile(conn, (1, 8)).ravel()
    A = sp.coo_matrix((K.ravel(), (rows, cols)), shape=(m ** 3, m ** 3)).tocsr()
    inner = np.zeros((m, m, m), dtype=bool)
    inner[1:-1, 1:-1, 1:-1] = True
    keep = np.flatnonzero(inner.ravel())
    A = A[keep][:, keep].tocsr()
    A = A / h  # O(1) entries
    A.sort_indices()
    b = A @ np.ones(A.shape[0])
    return A, b


def momentum3(n, seed=20261, peclet=2.0, coupling=0.05, time_term=0.5):
    u0 = np.array([1.0, 0.5, 0.25])
    vel = u0[None, :] + 0.1 * uniform_pm1(seed, np.arange(3 * n ** 3)).reshape(n ** 3, 3)
    nu = np.linalg.norm(u0) / peclet  # h = 1
    idx = np.arange(n ** 3)
    i, j, k = idx % n, (idx // n) % n, idx // (n * n)
    coord = (i, j, k)
    stride = (1, n, n * n)
    rows, cols, vals = [], [], []
    diag = np.full(n ** 3, time_term)
    for d in range(3):
        u = vel[:, d]
        lo = nu + np.maximum(u, 0.0)  # coefficient magnitude on the (i - 1) neighbour (upwind for u > 0)
        hi = nu + np.maximum(-u, 0.0)
        diag += lo + hi
        has_lo = coord[d] > 0
        has_hi = coord[d] < n - 1
        rows += [idx[has_lo], idx[has_hi]]
        cols += [idx[has_lo] - stride[d], idx[has_hi] + stride[d]]
        vals += [-lo[has_lo], -hi[has_hi]]
    rows.append(idx); cols.append(idx); vals.append(diag)
    S = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(n ** 3, n ** 3)).tocsr()
    M = np.array([[1.0, coupling, -coupling], [-coupling, 1.0, coupling], [coupling, -coupling, 1.0]])
    A = sp.kron(S, sp.csr_matrix(M), format="csr")  # node-blocked: row = 3 * node + dof
    A.sort_indices()
    b = A @ np.ones(A.shape[0])
    return A, b


def knn_graph_matrix(rng, n, dim=3, k=8, hubs=0, hub_deg=150, positive_frac=0.0, shift=0.01):
    """Unstructured test operator: weighted k-nearest-neighbour graph of n random points in the unit square / cube, symmetrised;
    `hubs` rows get `hub_deg` extra random connections (rows of hundreds of entries: the capacity tiers and overflow paths of the
    interpolation and SpGEMM kernels); a fraction of the off-diagonal pairs is made positive (the sign rules of the interpolation
    formulas); the diagonal is (1 + shift) times the absolute row sum.  Symmetric, diagonally dominant; rhs = A 1 has solution 1."""
    from scipy.spatial import cKDTree
    pts = rng.random((n, dim))
    d, idx = cKDTree(pts).query(pts, k=k + 1)
    rows = np.repeat(np.arange(n), k)
    cols = idx[:, 1:].ravel()
    w = 1.0 / (d[:, 1:].ravel() + 0.05)
    for h in rng.choice(n, size=hubs, replace=False):
        nb = rng.choice(n, size=hub_deg, replace=False)
        nb = nb[nb != h]
        rows = np.concatenate([rows, np.full(len(nb), h)])
        cols = np.concatenate([cols, nb])
        w = np.concatenate([w, rng.random(len(nb)) + 0.1])
    W = sp.coo_matrix((w, (rows, cols)), shape=(n, n)).tocsr()
    W = W.maximum(W.T).tocoo()
    sign = np.ones(W.nnz)
    if positive_frac > 0:
        up = W.row < W.col
        flip = up & (rng.random(W.nnz) < positive_frac)
        key = set(zip(W.row[flip].tolist(), W.col[flip].tolist()))
        sign = np.array([-1.0 if ((r, c) in key or (c, r) in key) else 1.0 for r, c in zip(W.row.tolist(), W.col.tolist())])
    off = sp.coo_matrix((-W.data * sign, (W.row, W.col)), shape=(n, n)).tocsr()
    diag = np.asarray(abs(off).sum(axis=1)).ravel() * (1.0 + shift) + 1e-3
    A = (off + sp.diags(diag)).tocsr()
    A.sort_indices()
    return A


def write_matrix_market(A, b, mat_path, rhs_path):
    """MatrixMarket coordinate real general matrix + array rhs, 1-based"""
    A = A.tocoo()
    with open(mat_path, "w") as f:
        f.write("%%MatrixMarket matrix coordinate real general\n")
        f.write(f"{A.shape[0]} {A.shape[1]} {A.nnz}\n")
        for r, c, v in zip(A.row, A.col, A.data):
            f.write(f"{r + 1} {c + 1} {v:.17g}\n")
    with open(rhs_path, "w") as f:
        f.write("%%MatrixMarket matrix array real general\n")
        f.write(f"{len(b)} 1\n")
        for v in b:
            f.write(f"{v:.17g}\n")


def write_hypre_ij(A, b, mat_prefix, rhs_prefix, nparts=1):
    """HYPRE IJ text files `<prefix>.<part 5 digits>`, 0-based ids (HypreSystem.cpp:1138-1318 reads these)"""
    A = A.tocsr()
    n = A.shape[0]
    q, r = divmod(n, nparts)
    for p in range(nparts):
        lo = q * p + min(p, r)
        hi = q * (p + 1) + min(p + 1, r) - 1
        with open(f"{mat_prefix}.{p:05d}", "w") as f:
            f.write(f"{lo} {hi} {lo} {hi}\n")
            for i in range(lo, hi + 1):
                for kk in range(A.indptr[i], A.indptr[i + 1]):
                    f.write(f"{i} {A.indices[kk]} {A.data[kk]:.17g}\n")
        with open(f"{rhs_prefix}.{p:05d}", "w") as f:
            f.write(f"{lo} {hi}\n")
            for i in range(lo, hi + 1):
                f.write(f"{i} {b[i]:.17g}\n")
