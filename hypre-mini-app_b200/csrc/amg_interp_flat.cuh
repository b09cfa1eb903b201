// amg_interp_flat.cuh -- extended+i interpolation rows, "flattened item" formulation (included by amg_interp.cu only).
// Same contract as rows::interp_row (amg_rows.cuh): bit-identical C-hat order, weights and truncation.
// Replaces hypre_BoomerAMGBuildExtPIInterp + hypre_BoomerAMGInterpTruncation reached from HYPRE_BoomerAMGSetup
// (/root/reference/src/HypreSystem.cpp:197,212,232 options, :323 via :692 call).
//
// G lanes own one row (G = 8, 16 or 32, chosen per level by the tier driver); BT / G rows per block of BT = 128 threads.  What changed
// against the first kernel (profiles/r01_interp_sass_notes.md:
// 6 078 warp instructions per row on level 1, control-flow bound):
//   * the walks run over C-COMPRESSED rows.  Of the row of a distributed neighbour j only the entries whose column is a C
//     point can be in C-hat_i, plus the one back-edge a_ji; of its strength row only the C entries can enter C-hat_i.  Those
//     are 14-31 % of the entries, so the tier driver builds, once per level and for F rows only, copies of A and S restricted
//     to C columns (CComp: original entry positions kept, they define the summation order) and the back-edge is found by a
//     binary search per neighbour (own rows are sorted by column after the diagonal; fetched halo rows are scanned);
//   * no per-neighbour warp loop.  The work of one row is a list of ITEMS -- (strong neighbour t, entry s of its strength
//     row) in the discovery, (distributed neighbour t, entry s of its matrix row) in the weights -- and the lanes take G
//     consecutive items per round, so every round is a coalesced load + one hash probe at full lane occupancy;
//   * the discovery order is the item order, so new C points take their list positions from a ballot prefix; an item that
//     repeats inside one round is removed with match.any;
//   * strong F neighbours of i are put into the same hash table with a MARK position: "is this entry of row i strong" is
//     the probe the weights phase does anyway (the first kernel re-shuffled the whole strength row for every chunk);
//   * contributions are staged neighbour-major; the denominator of a neighbour is summed in storage order by ONE lane per
//     neighbour (all neighbours of the chunk at once, one division per lane instead of one per warp step);
//   * every C-hat slot keeps a bit mask of the neighbours (tags) that contribute to it; a lane per slot then adds its
//     terms in ascending tag order, which is the order of the sequential loop over the entries of row i;
//   * the diagonal accumulates its at most one term per tag through a shuffle chain in tag order;
//   * truncation (qsort2abs tie order, sequential by definition) is a kernel of its own, k_interp_trunc, with one THREAD per row:
//     rows that need it are written raw to a chunked scratch and marked pending (MODE 2), so the row kernel has no block barrier
//     (one lane of every group, and later one warp per block behind a barrier, both left the machine idle);
//   * every collective names the full warp with width = G and every loop that contains one has a warp-uniform trip count
//     (per-group masks make the compiler serialise collectives over the groups, profiles/r01_interp_sass_notes.md).
#pragma once

namespace flat {

constexpr int MARK = 255;      // hash position of a strong F neighbour of i (not a C-hat slot)
constexpr unsigned FULL = 0xffffffffu;
// threads per block of the row kernel.  128 (8 resident blocks per SM) instead of 256 (4): a block keeps its shared memory until
// its slowest warp is done while the warps of C points leave at once, and the smaller the block the shorter that tail
// (ncu r02f: 36 % of the warp slots occupied at a limit of 50 % with 256 threads; measured: interp 48.2 -> 45.7 ms; 64 threads: no
// further gain)
constexpr int BT = 128;

// C-compressed copies of the level matrix and of the strength pattern (rows of non-isolated F points only, other rows empty):
// entry e of row j of AC is the entry acpos[e] of A (acpos ascending inside a row), whose column accol[e] is a C point.
struct CComp {
  const int *acptr, *accol, *acpos;
  const double *acval;
  const int *scptr, *sccol;
};

template <int G, int H, int CAP, int STG>
struct alignas(8) Grp {
  double lv[CAP];
  double stg_a[STG];
  double tval[G];
  int keys[H];
  int lc[CAP];
  unsigned smask[CAP];
  int noff[G + 1];
  int tb[G];
  int trb[G];
  int tcnt[G];
  unsigned short tseg[G];
  unsigned short stg_k[STG];   // position of the staged entry inside its row of A (rows of 65535+ entries take the fallback)
  int L, nout, row, flags;     // flags: 1 = any_off, 2 = F row with weights computed (truncate + write), 4 = write len only
  unsigned char spos[H];
  unsigned char stg_slot[STG];
  unsigned char tkind[G];
  unsigned char tneg[G];
  int pad_odd;                 // keeps sizeof / 4 odd: lane = group accesses of the truncation warp spread over the banks
};

template <int H>
__device__ __forceinline__ unsigned slot0(int key) {
  return (((unsigned)key * 2654435761u) >> 7) & (unsigned)(H - 1);
}
// -1: absent; 0 .. CAP-1: C-hat position; MARK: strong F neighbour of i
template <int H>
__device__ __forceinline__ int probe(const int *keys, const unsigned char *spos, int key) {
  unsigned h = slot0<H>(key);
  while (true) {
    const int k = keys[h];
    if (k == key) return spos[h];
    if (k == -1) return -1;
    h = (h + 1) & (unsigned)(H - 1);
  }
}
template <int H>
__device__ __forceinline__ void insert(int *keys, unsigned char *spos, int key, int pos) {  // key is known to be absent
  unsigned h = slot0<H>(key);
  while (true) {
    const int old = atomicCAS(keys + h, -1, key);
    if (old == -1) { spos[h] = (unsigned char)pos; return; }
    h = (h + 1) & (unsigned)(H - 1);
  }
}
template <int H>
__device__ __forceinline__ void set_pos(const int *keys, unsigned char *spos, int key, int pos) {  // key is present
  unsigned h = slot0<H>(key);
  while (keys[h] != key) h = (h + 1) & (unsigned)(H - 1);
  spos[h] = (unsigned char)pos;
}

template <int G>
__device__ __forceinline__ unsigned gballot(bool pred, int gbase) {
  const unsigned b = __ballot_sync(FULL, pred);
  if (G >= 32) return b;
  return (b >> gbase) & ((1u << (G & 31)) - 1u);
}
// tag of an item: the last t with noff[t] <= item (noff[0 .. G] exclusive offsets, non-decreasing, item < noff[G]).  Binary
// search: log2 G shared-memory reads, whatever the number of empty tags in between
template <int G>
__device__ __forceinline__ int tag_of(const int *noff, int item) {
  int t = 0;
#pragma unroll
  for (int step = G >> 1; step > 0; step >>= 1)
    if (noff[t + step] <= item) t += step;
  return t;
}
// inclusive scan over the G lanes of a group (full-warp shuffle, width G)
template <int G>
__device__ __forceinline__ int gscan_incl(int v, int gl) {
#pragma unroll
  for (int d = 1; d < G; d <<= 1) {
    const int u = __shfl_up_sync(FULL, v, d, G);
    if (gl >= d) v += u;
  }
  return v;
}

// lraw / len / MODE / lo / row_list: as for the tier driver in amg_interp.cu (MODE 0 count, 1 fill at Pptr, 2 single pass at
// i * pmax).  Scf[k] = cf[Scol[k]] as a signed char.
template <int G, int H, int CAP, int STG, int MODE>
__global__ void __launch_bounds__(BT, 1024 / BT)
k_interp_flat(rows::Csr A, const int *__restrict__ Sptr, const int *__restrict__ Scol, const signed char *__restrict__ Scf,
              CComp cc, const int *__restrict__ cf, const int *__restrict__ f2c, int n_own, int n_rows, double tf, int pmax, int lo, int *lraw,
              int *len, const int *__restrict__ Pptr, int *Pcol, double *Pval, const int *__restrict__ row_list,
              int slot0, int *scr_c, double *scr_v, unsigned long long *tstats) {
  using GS = Grp<G, H, CAP, STG>;
  // tstats != nullptr (B200_INTERP_TIMING=1): cycles per phase summed over the warps of the launch -- [0] discovery,
  // [1] weights, [2] wait at the first block barrier, [3] truncation (first warp), [4] wait at the second barrier, [5] write
  long long tmark = tstats ? clock64() : 0;
  auto lap = [&](int slot) {
    if (tstats && (threadIdx.x & 31) == 0) {
      const long long now = clock64();
      atomicAdd(tstats + slot, (unsigned long long)(now - tmark));
      tmark = now;
    }
  };
  static_assert(CAP <= 128 && G <= 32 && STG > CAP, "slot / tag encodings");
  // 1024 resident threads per SM: 227 KB of shared memory, 1 KB per block reserved by the system.  The kernel is latency bound,
  // occupancy is what hides its dependent loads (r02c: a footprint one block too large cost the whole gain of that step)
  static_assert((sizeof(GS) * (BT / G) + 1024) * (1024 / BT) <= 227 * 1024, "shared memory of the resident blocks exceeds the SM");
  constexpr int GROUPS = BT / G;
  constexpr int HMAX = (H * 3) / 4;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  GS *all = reinterpret_cast<GS *>(smem_raw);
  const int g = threadIdx.x / G, gl = threadIdx.x % G;
  const int lane32 = threadIdx.x & 31;
  const int gbase = lane32 - gl;
  const unsigned ltm = (1u << gl) - 1u;
  GS &S = all[g];
  // MODE 2 defers truncation: a row that needs it is written raw (C-hat list + weights) at (slot - slot0) * CAP of the scratch
  // arrays and finished by k_interp_trunc (one THREAD per row); this kernel then has no block barrier at all.
  constexpr bool DEFER = (MODE == 2);
  const int slot = slot0 + blockIdx.x * GROUPS + g;
  int i = -1;
  if (slot < n_rows) i = row_list ? row_list[slot] : slot;

  // ---- what is there to do for this row ---------------------------------------------------------------------------------
  // all header words of the row are requested together (one round trip to memory instead of four dependent ones)
  bool active = false;
  int h_lraw = 0, h_cf = 1, h_sb = 0, h_se = 0, h_ab = 0, h_ae = 0;
  if (i >= 0) {
    h_lraw = lraw[i];
    h_cf = cf[i];
    h_sb = Sptr[i];
    h_se = Sptr[i + 1];
    h_ab = A.ptr[i];
    h_ae = A.ptr[i + 1];
  }
  // the first chunk of row i and the extents of its neighbour rows do not depend on the discovery: their loads are issued
  // now and are in flight while C-hat is built (classification after the discovery only adds the hash probe)
  int p_i1 = -1, p_rb = 0, p_re = 0, p_b1 = 0, p_nt = 0;
  double p_aij = 0.0;
  {
    const int jj = h_ab + 1 + gl;
    if (i >= 0 && jj < h_ae) {
      p_i1 = A.col[jj];
      p_aij = A.val[jj];
      p_rb = A.ptr[p_i1];
      p_re = A.ptr[p_i1 + 1];
      p_b1 = cc.acptr[p_i1];
      p_nt = cc.acptr[p_i1 + 1] - p_b1;
    }
  }
  if (i >= 0) {
    // MODE 0 / 2: rows not computed yet (lraw == -1).  MODE 1 (fill): rows counted but not filled yet (lraw >= 0) that fit this
    // tier; a filled row is marked lraw = -2 - L, so a row that this tier has to give up (hash load) is still taken by the next
    bool todo;
    if (MODE != 1) todo = (h_lraw == -1);
    else todo = (h_lraw >= 0 && h_lraw <= CAP);
    if (todo) {
      const int cfi = h_cf;
      if (cfi >= 0) {   // C point: identity row
        if (gl == 0) {
          if (MODE != 1) { lraw[i] = 0; len[i] = 1; }
          if (MODE != 0) { const size_t b = (MODE == 1) ? (size_t)Pptr[i] : (size_t)i * pmax; Pcol[b] = f2c[i]; Pval[b] = 1.0; }
          if (MODE == 1) lraw[i] = -2;
        }
      } else if (cfi == -3) {
        if (gl == 0) { if (MODE != 1) { lraw[i] = 0; len[i] = 0; } else lraw[i] = -2; }
      } else if (MODE == 1 && Pptr[i + 1] == Pptr[i]) {
        if (gl == 0) lraw[i] = -2;
      } else active = true;
    }
  }
  for (int s = gl; s < H; s += G) S.keys[s] = -1;
  if (gl == 0) { S.L = 0; S.nout = 0; S.row = i; S.flags = 0; }
  __syncwarp(FULL);

  // ---- discovery of C-hat_i: items in sequence order ---------------------------------------------------------------------
  int L = 0, nmark = 0;
  bool ovf = false, any_off = false;
  {
    const int sb = active ? h_sb : 0, se = active ? h_se : 0;
    const int nchunk = __reduce_max_sync(FULL, (se - sb + G - 1) / G);
    for (int c = 0; c < nchunk; ++c) {
      const int kk = sb + c * G + gl;
      const bool valid = (kk < se) && !ovf;
      int i1 = -1, c1 = -3;
      if (valid) { i1 = Scol[kk]; c1 = Scf[kk]; }
      const bool isC = valid && c1 >= 0;
      const bool isF = valid && c1 < 0 && c1 != -3;
      int b2 = 0, n_t = 0;
      if (isC) { b2 = ~i1; n_t = 1; }                                 // the item is the neighbour itself
      else if (isF) { b2 = cc.scptr[i1]; n_t = cc.scptr[i1 + 1] - b2; }   // the items are the C entries of its strength row
      const unsigned fmk = gballot<G>(isF, gbase);
      if (L + nmark + __popc(fmk) > HMAX) ovf = true;
      if (isF && !ovf) insert<H>(S.keys, S.spos, i1, MARK);
      nmark += __popc(fmk);
      const int incl = gscan_incl<G>(ovf ? 0 : n_t, gl);
      const int T = __shfl_sync(FULL, incl, G - 1, G);
      S.noff[gl] = incl - (ovf ? 0 : n_t);
      if (gl == G - 1) S.noff[G] = T;
      S.tb[gl] = b2;
      __syncwarp(FULL);
      const int Tmax = __reduce_max_sync(FULL, T);
      for (int base = 0; base < Tmax; base += G) {
        const int item = base + gl;
        int key = -1;
        bool cand = false;
        if (item < T && !ovf) {
          const int tp = tag_of<G>(S.noff, item);
          const int bb = S.tb[tp];
          if (bb < 0) { key = ~bb; cand = true; }
          else { key = cc.sccol[bb + (item - S.noff[tp])]; cand = true; }
        }
        const bool nw = cand && (probe<H>(S.keys, S.spos, key) < 0);
        const unsigned long long mkey = nw ? (((unsigned long long)(unsigned)g << 32) | (unsigned)key)
                                           : (0xffffffff00000000ull | (unsigned)lane32);
        const unsigned same = __match_any_sync(FULL, mkey);
        const bool first = nw && ((__ffs(same) - 1) == lane32);
        const unsigned nm = gballot<G>(first, gbase);
        const unsigned offm = gballot<G>(first && key >= n_own, gbase);
        const int cnt = __popc(nm);
        if (cnt) {
          if (L + cnt > CAP || L + cnt + nmark > HMAX) ovf = true;
          else {
            if (first) {
              const int pos = L + __popc(nm & ltm);
              insert<H>(S.keys, S.spos, key, pos);
              S.lc[pos] = key;
            }
            any_off |= (offm != 0u);
            L += cnt;
          }
        }
        __syncwarp(FULL);
      }
      __syncwarp(FULL);
    }
  }
  if (ovf) active = false;   // lraw stays -1: the next tier takes the row
  lap(0);
  const bool full = (MODE != 0) || tf != 0.0;
  if (!full) {   // uniform over the grid
    if (active && gl == 0) { lraw[i] = L; len[i] = (pmax > 0 && L > pmax) ? pmax : L; }
    return;
  }

  // ---- own-range C points first (P_diag before P_offd), stable: positions change, the table follows ---------------------
  if (__any_sync(FULL, active && any_off)) {
    const bool act = active && any_off;
    int *tmp = reinterpret_cast<int *>(S.smask);
    const int Lmax = __reduce_max_sync(FULL, act ? L : 0);
    int nown = 0;
    for (int base = 0; base < Lmax; base += G) {
      const int k = base + gl;
      const bool own = act && (k < L) && (S.lc[k] < n_own);
      nown += __popc(gballot<G>(own, gbase));
    }
    int co = 0, cx = 0;
    for (int base = 0; base < Lmax; base += G) {
      const int k = base + gl;
      const bool valid = act && k < L;
      const int v = valid ? S.lc[k] : 0;
      const bool own = valid && v < n_own, off = valid && !own;
      const unsigned mo = gballot<G>(own, gbase), mx = gballot<G>(off, gbase);
      if (valid) {
        const int np = own ? (co + __popc(mo & ltm)) : (nown + cx + __popc(mx & ltm));
        tmp[np] = v;
        set_pos<H>(S.keys, S.spos, v, np);
      }
      co += __popc(mo);
      cx += __popc(mx);
    }
    __syncwarp(FULL);
    if (act)
      for (int k = gl; k < L; k += G) S.lc[k] = tmp[k];
    __syncwarp(FULL);
  }

  // ---- weights ------------------------------------------------------------------------------------------------------------
  for (int k = gl; k < CAP; k += G) S.lv[k] = 0.0;
  const int ab = active ? h_ab : 0, ae = active ? h_ae : 0;
  double diagonal = active ? A.val[ab] : 1.0;
  const int na = active ? (ae - ab - 1) : 0;
  const int nchunkA = __reduce_max_sync(FULL, (na + G - 1) / G);
  __syncwarp(FULL);
  for (int c = 0; c < nchunkA; ++c) {
    // one entry of row i per lane (tag = lane): classification
    const int jj = ab + 1 + c * G + gl;
    const bool valid = active && jj < ae;
    int kind = 3, pr = -1, b1 = 0, n_t = 0, kb = -1, rb0 = 0;   // kb: position of the back-edge a_{i1,i} inside row i1 (-1: none / wrong sign)
    bool neg = false, long_row = false;
    double aij = 0.0, ab_val = 0.0;
    if (valid) {
      int i1, rb, re;
      if (c == 0) { i1 = p_i1; aij = p_aij; rb = p_rb; re = p_re; b1 = p_b1; n_t = p_nt; }
      else {
        i1 = A.col[jj];
        aij = A.val[jj];
        rb = A.ptr[i1];
        re = A.ptr[i1 + 1];
        b1 = cc.acptr[i1];
        n_t = cc.acptr[i1 + 1] - b1;
      }
      pr = probe<H>(S.keys, S.spos, i1);
      if (pr >= 0 && pr != MARK) kind = 0;                    // direct: a_ij goes to its own C point
      else if (pr == MARK) {                                   // strong F neighbour: distributed over C-hat and i
        kind = 1;
        rb0 = rb;
        neg = A.val[rb] < 0;
        // back-edge a_{i1,i}: entries after the diagonal are sorted by column in own rows, in storage order of the owner in
        // fetched halo rows
        if (i1 < n_own) {
          int lo_k = rb + 1, hi_k = re;
          while (lo_k < hi_k) {
            const int mid = lo_k + ((hi_k - lo_k) >> 1);   // entry positions reach 2^31: no lo + hi
            if (A.col[mid] < i) lo_k = mid + 1; else hi_k = mid;
          }
          if (lo_k < re && A.col[lo_k] == i) kb = lo_k;
        } else {
          for (int k = rb + 1; k < re; ++k)
            if (A.col[k] == i) { kb = k; break; }
        }
        if (kb >= 0) {
          ab_val = A.val[kb];
          kb -= rb;                                            // position inside the row, like the staged entries
          if (!(neg ? (ab_val > 0) : (ab_val < 0))) kb = -1;   // wrong sign: does not take part
        }
        if (re - rb >= 65535) long_row = true;
      } else if (cf[i1] != -3) kind = 2;                       // lumped into the diagonal
    }
    const int ntag = active ? max(0, min(G, na - c * G)) : 0;
    S.tb[gl] = b1;
    S.trb[gl] = rb0;
    S.tneg[gl] = neg ? 1 : 0;
    if (__any_sync(FULL, long_row)) {   // 16-bit positions do not reach: the group gives the row to the fallback path
      if (gballot<G>(long_row, gbase)) { active = false; ovf = true; }
    }
    S.tkind[gl] = (unsigned char)kind;
    int lo_t = 0, rlen = G;   // tag range of the current attempt: [lo_t, lo_t + rlen)
    while (__any_sync(FULL, active && lo_t < ntag)) {
      const bool run = active && lo_t < ntag;
      const bool in = run && gl >= lo_t && gl < lo_t + rlen && gl < ntag;
      const int n_in = (in && kind == 1) ? n_t : 0;
      const int incl = gscan_incl<G>(n_in, gl);
      const int T = __shfl_sync(FULL, incl, G - 1, G);
      S.noff[gl] = incl - n_in;
      if (gl == G - 1) S.noff[G] = T;
      S.tcnt[gl] = 0;
      for (int k = gl; k < L; k += G) S.smask[k] = 0u;
      __syncwarp(FULL);
      if (in && kind == 0) atomicOr(&S.smask[pr], 1u << gl);
      // walk: G consecutive items per round
      const int Tmax = __reduce_max_sync(FULL, T);
      int nstg = 0;
      bool sovf = false;
      for (int base = 0; base < Tmax; base += G) {
        const int item = base + gl;
        bool contrib = false;
        int sl = 0;
        double a = 0.0;
        int kpos = 0, tp = 0;
        if (item < T) {
          tp = tag_of<G>(S.noff, item);
          const int e1 = S.tb[tp] + (item - S.noff[tp]);
          a = cc.acval[e1];
          const bool ok = S.tneg[tp] ? (a > 0) : (a < 0);
          if (ok) {
            const int p2 = probe<H>(S.keys, S.spos, cc.accol[e1]);
            if (p2 >= 0 && p2 != MARK) { contrib = true; sl = p2; kpos = cc.acpos[e1] - S.trb[tp]; }
          }
        }
        const unsigned cm = gballot<G>(contrib, gbase);
        const int cn = __popc(cm);
        if (nstg + cn > STG) sovf = true;
        if (contrib && !sovf) {
          const int e = nstg + __popc(cm & ltm);
          S.stg_slot[e] = (unsigned char)sl;
          S.stg_a[e] = a;
          S.stg_k[e] = (unsigned short)kpos;
          atomicAdd(&S.tcnt[tp], 1);
          atomicOr(&S.smask[sl], 1u << tp);
        }
        nstg += cn;
      }
      __syncwarp(FULL);
      // per tag: denominator in storage order, distribute factor, diagonal term
      const bool commit = run && !sovf;
      const int cnt_t = S.tcnt[gl];
      const int seg = gscan_incl<G>(cnt_t, gl) - cnt_t;
      double tv = 0.0, dterm = 0.0;
      bool hasd = false;
      if (commit && in) {
        if (kind == 1) {
          // denominator: contributing entries of the neighbour row in storage order; the back-edge joins at its position
          double sum = 0.0;
          bool pending = (kb >= 0);
          for (int e = seg; e < seg + cnt_t; ++e) {
            if (pending && kb < (int)S.stg_k[e]) { sum = sum + ab_val; pending = false; }
            sum = sum + S.stg_a[e];
          }
          if (pending) sum = sum + ab_val;
          if (sum != 0.0) {
            tv = aij / sum;
            if (kb >= 0) { dterm = tv * ab_val; hasd = true; }
          } else { dterm = aij; hasd = true; }
        } else if (kind == 0) tv = aij;
        else if (kind == 2) { dterm = aij; hasd = true; }
      }
      S.tval[gl] = tv;
      S.tseg[gl] = (unsigned short)seg;
      __syncwarp(FULL);
      // per C-hat slot: terms in ascending tag order
      if (commit) {
        for (int p = gl; p < L; p += G) {
          unsigned m = S.smask[p];
          double acc = S.lv[p];
          while (m) {
            const int t = __ffs(m) - 1;
            m &= m - 1;
            double term;
            if (S.tkind[t] == 1) {
              int e = S.tseg[t];
              while (S.stg_slot[e] != p) ++e;
              term = S.tval[t] * S.stg_a[e];
            } else term = S.tval[t];
            acc = acc + term;
          }
          S.lv[p] = acc;
        }
      }
      // diagonal: at most one term per tag, in tag order; every lane of the group keeps the same running value
      unsigned hm = gballot<G>(hasd, gbase);
      __syncwarp(FULL);          // the slot pass has read tval: reuse it for the diagonal terms
      S.tval[gl] = dterm;
      __syncwarp(FULL);
      while (hm) {
        const int t = __ffs(hm) - 1;
        hm &= hm - 1;
        diagonal = diagonal + S.tval[t];
      }
      if (run) {
        if (!sovf) lo_t += rlen;
        else if (rlen == 1) { active = false; ovf = true; }   // one neighbour alone exceeds the staging area: next tier
        else rlen >>= 1;
      }
      __syncwarp(FULL);
    }
  }
  if (active && diagonal != 0.0) {
    const double md = -diagonal;
    for (int k = gl; k < L; k += G) S.lv[k] = S.lv[k] / md;
  }
  lap(1);
  if (DEFER) {
    if (!active) return;
    const bool needs_trunc = (tf > 0.0 && L > 0) || (pmax > 0 && L > pmax);
    if (needs_trunc) {
      const size_t sb0 = (size_t)(slot - slot0) * CAP;
      for (int k = gl; k < L; k += G) { scr_c[sb0 + k] = S.lc[k]; scr_v[sb0 + k] = S.lv[k]; }
      if (gl == 0) { lraw[i] = L; len[i] = any_off ? -2 : -1; }   // pending: k_interp_trunc sets the final length
    } else {
      if (gl == 0) { lraw[i] = L; len[i] = L; }
      const size_t b = (size_t)i * pmax;
      for (int k = gl; k < L; k += G) { Pcol[b + k] = f2c[S.lc[k]]; Pval[b + k] = S.lv[k]; }
    }
    lap(5);
    return;
  }
  if (gl == 0) { S.L = L; S.nout = L; S.flags = active ? (2 | (any_off ? 1 : 0)) : 0; }
  __syncthreads();
  lap(2);

  // ---- truncation: sequential per row (the sums and hypre's qsort2abs are order dependent); lane = row of the block -------
  if (threadIdx.x < GROUPS) {
    GS &T = all[threadIdx.x];
    if (T.flags & 2) {
      int *lc = T.lc;
      double *lv = T.lv;
      int n_out = T.L;
      if (tf > 0.0 && n_out > 0) {
        double max_coef = 0.0, row_sum = 0.0, scale = 0.0;
        for (int k = 0; k < n_out; ++k) { const double a = fabs(lv[k]); max_coef = (a > max_coef) ? a : max_coef; }
        max_coef = max_coef * tf;
        int w = 0;
        for (int k = 0; k < n_out; ++k) {
          row_sum = row_sum + lv[k];
          if (fabs(lv[k]) >= max_coef) { scale = scale + lv[k]; lc[w] = lc[k]; lv[w] = lv[k]; ++w; }
        }
        n_out = w;
        if (scale != 0.0 && scale != row_sum) {
          const double f = row_sum / scale;
          for (int k = 0; k < n_out; ++k) lv[k] = lv[k] * f;
        }
      }
      if (pmax > 0 && n_out > pmax) {
        double row_sum = 0.0, scale = 0.0;
        for (int k = 0; k < n_out; ++k) row_sum = row_sum + lv[k];
        rows::qsort2abs(lc, lv, n_out, pmax);
        for (int k = 0; k < pmax; ++k) scale = scale + lv[k];
        n_out = pmax;
        if (scale != 0.0 && scale != row_sum) {
          const double f = row_sum / scale;
          for (int k = 0; k < n_out; ++k) lv[k] = lv[k] * f;
        }
        if (T.flags & 1) {   // re-split, each part keeping the sorted order
          for (int k = 1; k < n_out; ++k) {
            const int v = lc[k];
            const double x = lv[k];
            if (v < n_own) {
              int p = k;
              while (p > 0 && lc[p - 1] >= n_own) { lc[p] = lc[p - 1]; lv[p] = lv[p - 1]; --p; }
              lc[p] = v;
              lv[p] = x;
            }
          }
        }
      }
      T.nout = n_out;
    }
  }
  lap(3);
  __syncthreads();
  lap(4);

  // ---- write ---------------------------------------------------------------------------------------------------------------
  if (!(S.flags & 2)) return;
  const int n_out = S.nout;
  if (MODE != 1 && gl == 0) { lraw[i] = L; len[i] = n_out; }
  if (MODE == 1 && gl == 0) lraw[i] = -2 - L;
  if (MODE == 0) return;
  const size_t b = (MODE == 1) ? (size_t)Pptr[i] : (size_t)i * pmax;
  for (int k = gl; k < n_out; k += G) {
    Pcol[b + k] = f2c[S.lc[k]];
    Pval[b + k] = S.lv[k];
  }
  lap(5);
}

// Deferred truncation (MODE 2): one thread per row on the raw lists the row kernel left in the scratch arrays.  The sums and
// hypre's qsort2abs are sequential per row by definition; here 32 rows advance per warp instruction, instead of one lane of a
// group (or one warp of a block) working alone.  Lists of up to TRUNC_STAGE entries (nearly all) are sorted in shared memory
// (row stride TRUNC_STAGE + 1: lanes that touch the same entry index hit different banks); longer ones in the global scratch.
// The first version worked on the global scratch only and was bound by dependent L1 / L2 loads (ncu: long scoreboard 64
// cycles per issue, profiles/r02b_trunc_ncu_summary.txt).
constexpr int TRUNC_STAGE = 32;
constexpr int TRUNC_THREADS = 96;    // 96 x 33 x 12 B = 38 KB of static shared memory, 5 blocks per SM
__global__ void __launch_bounds__(TRUNC_THREADS)
k_interp_trunc(int slot0, int slot_end, const int *__restrict__ row_list, int cap, int *scr_c, double *scr_v, const int *__restrict__ lraw,
               int *len, const int *__restrict__ f2c, int n_own, double tf, int pmax, int *Pcol, double *Pval) {
  __shared__ double s_v[TRUNC_THREADS][TRUNC_STAGE + 1];
  __shared__ int s_c[TRUNC_THREADS][TRUNC_STAGE + 1];
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int slot = slot0 + t;
  if (slot >= slot_end) return;
  const int i = row_list ? row_list[slot] : slot;
  const int pend = len[i];
  if (pend >= 0) return;
  const bool any_off = (pend == -2);
  int n_out = lraw[i];
  int *lc = scr_c + (size_t)t * cap;
  double *lv = scr_v + (size_t)t * cap;
  if (n_out <= TRUNC_STAGE) {
    int *sc = s_c[threadIdx.x];
    double *sv = s_v[threadIdx.x];
    for (int k = 0; k < n_out; ++k) { sc[k] = lc[k]; sv[k] = lv[k]; }
    lc = sc;
    lv = sv;
  }
  if (tf > 0.0 && n_out > 0) {
    double max_coef = 0.0, row_sum = 0.0, scale = 0.0;
    for (int k = 0; k < n_out; ++k) { const double a = fabs(lv[k]); max_coef = (a > max_coef) ? a : max_coef; }
    max_coef = max_coef * tf;
    int w = 0;
    for (int k = 0; k < n_out; ++k) {
      const double x = lv[k];
      row_sum = row_sum + x;
      if (fabs(x) >= max_coef) { scale = scale + x; lc[w] = lc[k]; lv[w] = x; ++w; }
    }
    n_out = w;
    if (scale != 0.0 && scale != row_sum) {
      const double f = row_sum / scale;
      for (int k = 0; k < n_out; ++k) lv[k] = lv[k] * f;
    }
  }
  if (pmax > 0 && n_out > pmax) {
    double row_sum = 0.0, scale = 0.0;
    for (int k = 0; k < n_out; ++k) row_sum = row_sum + lv[k];
    rows::qsort2abs(lc, lv, n_out, pmax);
    for (int k = 0; k < pmax; ++k) scale = scale + lv[k];
    n_out = pmax;
    if (scale != 0.0 && scale != row_sum) {
      const double f = row_sum / scale;
      for (int k = 0; k < n_out; ++k) lv[k] = lv[k] * f;
    }
    if (any_off) {   // re-split, each part keeping the sorted order
      for (int k = 1; k < n_out; ++k) {
        const int v = lc[k];
        const double x = lv[k];
        if (v < n_own) {
          int p = k;
          while (p > 0 && lc[p - 1] >= n_own) { lc[p] = lc[p - 1]; lv[p] = lv[p - 1]; --p; }
          lc[p] = v;
          lv[p] = x;
        }
      }
    }
  }
  len[i] = n_out;
  const size_t b = (size_t)i * pmax;
  for (int k = 0; k < n_out; ++k) { Pcol[b + k] = f2c[lc[k]]; Pval[b + k] = lv[k]; }
}

}  // namespace flat
