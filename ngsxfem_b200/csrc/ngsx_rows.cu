// ngsx_rows.cu -- (d') owner-computes assembly: plan construction (a static schedule) and the row-gather kernel.
//
// The coloured scatter (ngsx_assemble.cu) streams the whole CSR value array once per colour: on the 10M-tet config every
// 32-byte sector of the 584 MB value array is read and written by ~13 different launches.  Here the dependency is turned
// around: the plan lists, for every non-empty CSR row, the (element, local row) pairs that contribute to it, and ONE LANE
// gathers a row completely -- contributions are summed in plan order in the warp's shared-memory window and each value leaves
// the SM once, as a coalesced store.  No colouring, no zero-fill, no read-modify-write of global memory, no floating-point
// atomics; the summation order of every CSR entry is fixed by the plan, so the result is bitwise reproducible.
//
// Second generation (round 2).  The first one (per-row records, per-lane row ranges, reference tables in shared memory) ran at
// 26 % of the HBM peak by its real traffic: ncu showed the LSU / shared-memory pipe 86 % busy -- per pair two 32-byte loads of
// a 64-byte geometry record, a 16-byte descriptor load from a lane-private address, ~11 128-bit table reads and 20 scattered
// 8-byte window accesses.  Now:
//   * descriptors are stored in the order the warp consumes them (step-major, lane-minor, no padding): one coalesced 128-bit
//     load per step instead of 32 distinct lines, and no row records / split tables at all (ngsx_rows.cuh);
//   * rows inside a tile are dealt to the lanes by decreasing pair count (LPT), so the lanes of a tile finish together;
//   * the constant-rule row of an uncut element is computed WITHOUT tables: for the permutation-symmetric NGSolve rules the
//     row of the P1/P2 element matrix is a closed form in the rule's moments (kernel constants) and in the rows
//     K[a][.] = grad(lam_a) G grad(lam_.) of the element's geometry record, which k_geom stores per local vertex (32 bytes) --
//     a vertex row loads ONE 32-byte record, an edge row two, and no shared-memory table is read;
//   * cut elements read their row of the compact quadrature buffer as before.
#include <cub/cub.cuh>
#include <algorithm>
#include <numeric>
#include <vector>
#include <thrust/iterator/counting_iterator.h>
#include "ngsx_rows.cuh"
#include "rule_tables.h"

namespace ngsx {

// ---- plan construction --------------------------------------------------------------------------------------------
// key = (row, local row i, list index k): the pairs of a row are walked grouped by i (plan order = summation order)
__global__ void __launch_bounds__(256) k_pair_keys(int64_t npairs, int nd, int64_t nlist, const int32_t *__restrict__ table, int64_t row_lo,
                                                  int64_t row_hi, const int32_t *__restrict__ row_owner, int32_t me, uint64_t *__restrict__ keys) {
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npairs) return;
  int32_t r = table[p];
  const int64_t k = p / nd, i = p - k * nd;
  bool mine = r >= row_lo && r < row_hi;
  if (mine && row_owner) mine = row_owner[r] == me;
  keys[p] = mine ? (((uint64_t)(uint32_t)r << 32) | (uint64_t)(i * nlist + k)) : ~0ULL;
}

__global__ void k_lower_bound(int64_t n, const uint64_t *__restrict__ keys, uint64_t target, int64_t *out) {
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (keys[mid] < target) lo = mid + 1; else hi = mid;
  }
  *out = lo;
}

__global__ void __launch_bounds__(256) k_head_flags(int64_t n, const uint64_t *__restrict__ keys, uint8_t *__restrict__ flags) {
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  flags[p] = (p == 0) || ((keys[p] >> 32) != (keys[p - 1] >> 32));
}

// per plan row: CSR span {rowptr[r], rowptr[r+1]}; closes pairptr; row statistics
__global__ void __launch_bounds__(256) k_plan_rows(int64_t nrows, int64_t npairs, const uint64_t *__restrict__ keys, int64_t *__restrict__ pairptr,
                                                  const int64_t *__restrict__ rowptr, int64_t *__restrict__ span,
                                                  unsigned long long *__restrict__ stat /* {covered, maxlen, lo, hi} */) {
  int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q == 0) pairptr[nrows] = npairs;
  if (q >= nrows) return;
  int32_t r = (int32_t)(keys[pairptr[q]] >> 32);
  span[2 * q] = rowptr[r]; span[2 * q + 1] = rowptr[r + 1];
  unsigned long long len = (unsigned long long)(rowptr[r + 1] - rowptr[r]);
  atomicAdd(&stat[0], len);
  atomicMax(&stat[1], len);
  if (q == 0) stat[2] = (unsigned long long)rowptr[r];
  if (q == nrows - 1) stat[3] = (unsigned long long)rowptr[r + 1];
}

// schedule record of a plan row (host -> device, symbolic phase only)
struct RowSched { int32_t tile, start, lane, off; };

// one thread per pair: its slot in the schedule, the positions of the nd local columns inside the CSR row
__global__ void __launch_bounds__(128) k_plan_desc(int64_t npairs, int64_t nrows, int nd, int64_t nlist, int ds, const uint64_t *__restrict__ keys,
                                                  const int64_t *__restrict__ pairptr, const RowSched *__restrict__ sched,
                                                  const int64_t *__restrict__ tile_d0, const int64_t *__restrict__ tile_so,
                                                  const int32_t *__restrict__ stepoffs, const int32_t *__restrict__ table,
                                                  const int64_t *__restrict__ rowptr, const int32_t *__restrict__ colidx,
                                                  uint8_t *__restrict__ desc, int *__restrict__ bad, int lower) {
  const int64_t pp = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pp >= npairs) return;
  int64_t lo = 0, hi = nrows;                                   // plan row of the pair: last q with pairptr[q] <= pp
  while (hi - lo > 1) {
    const int64_t mid = (lo + hi) >> 1;
    if (pairptr[mid] <= pp) lo = mid; else hi = mid;
  }
  const int64_t q = lo;
  const RowSched rs = sched[q];
  const int step = rs.start + (int)(pp - pairptr[q]);
  const int64_t slot = tile_d0[rs.tile] + stepoffs[tile_so[rs.tile] + step] + rs.lane;
  const uint64_t key = keys[pp];
  const int32_t r = (int32_t)(key >> 32);
  const uint32_t p = (uint32_t)(key & 0xffffffffULL);
  const int i = (int)(p / (uint32_t)nlist);
  const uint32_t k = p - (uint32_t)i * (uint32_t)nlist;
  uint8_t *d = desc + slot * ds;
  *reinterpret_cast<uint32_t *>(d) = k | ((uint32_t)i << DESC_KBITS);
  *reinterpret_cast<uint16_t *>(d + 4) = (uint16_t)rs.off;
  const int64_t rs0 = rowptr[r], re = rowptr[r + 1];
  for (int j = 0; j < nd; j++) {
    const int32_t col = table[(int64_t)k * nd + j];
    int64_t a = rs0, b = re - 1;
    while (a < b) {
      const int64_t mid = (a + b) >> 1;
      if (colidx[mid] < col) a = mid + 1; else b = mid;
    }
    int64_t pos = a - rs0;
    if (lower && col > r) pos = POS_SKIP;                       // symmetric storage: the row keeps its columns <= row only
    else {
      if (b < rs0 || colidx[a] != col) { atomicOr(bad, 1); pos = 0; }
      if (pos >= (lower ? POS_SKIP : MAXROW)) { atomicOr(bad, 2); pos = 0; }
    }
    d[6 + j] = (uint8_t)pos;
  }
}

// processing order of the tiles: by the list index of the first pair of the tile, so that the vertex-row and edge-row tiles of
// one mesh region run at about the same time (their geometry records are then L2 hits)
// Small plans (the geometry records of the whole list fit in L2: a rank's share of a multi-GPU run) are ordered by decreasing
// step count instead, list index second: the warps claim tiles from a counter, so the longest tiles go first and the
// kernel's tail is made of short ones (at 1/8 of config 2 the SMs finished between 113 K and 190 K cycles in mesh order).
__global__ void __launch_bounds__(256) k_tile_keys(int64_t ntiles, const TileHdr *__restrict__ tiles, const uint8_t *__restrict__ desc, int ds,
                                                  int by_cost, uint32_t *__restrict__ keys, int32_t *__restrict__ ids) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ntiles) return;
  ids[c] = (int32_t)c;
  const uint32_t k = *reinterpret_cast<const uint32_t *>(desc + tiles[c].d0 * ds) & DESC_KMASK;
  keys[c] = by_cost ? ((uint32_t)(255 - min(tiles[c].nsteps, 255)) << 24) | (k >> 3) : k;
}
__global__ void __launch_bounds__(256) k_tile_permute(int64_t ntiles, const int32_t *__restrict__ order, const TileHdr *__restrict__ in,
                                                     TileHdr *__restrict__ out) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ntiles) return;
  out[c] = in[order[c]];
}

// slot of every (list entry, local row) pair -- what a producer of packed block rows needs (ngsx_rowplan_rowslots)
__global__ void __launch_bounds__(256) k_rowslots(int64_t nslots, int nd, int ds, const uint8_t *__restrict__ desc, int32_t *__restrict__ rowslot) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots) return;
  const uint32_t w = *reinterpret_cast<const uint32_t *>(desc + s * ds);
  rowslot[(int64_t)(w & DESC_KMASK) * nd + (w >> DESC_KBITS)] = (int32_t)s;
}

// ---- moments of the reference rule ----------------------------------------------------------------------------------
bool rule_moments(int dim, int order, RuleMom *out) {
  if (order < 0) order = 0;
  const int maxo = dim == 2 ? NGSX_RULE_TRIG_MAXORDER : NGSX_RULE_TET_MAXORDER;
  if ((dim != 2 && dim != 3) || order > maxo) return false;
  const int *offs = dim == 2 ? NGSX_RULE_TRIG_OFFS : NGSX_RULE_TET_OFFS;
  const double (*data)[4] = dim == 2 ? NGSX_RULE_TRIG_DATA : NGSX_RULE_TET_DATA;
  const int nv = dim + 1;
  double m1[4] = {0, 0, 0, 0}, m2[4][4] = {}, m3[4][4][4] = {}, m4[4][4][4][4] = {};
  double m0 = 0.0;
  for (int q = offs[order]; q < offs[order + 1]; q++) {
    double l[4] = {data[q][0], data[q][1], dim == 3 ? data[q][2] : 0.0, 0.0};
    double last = 1.0;
    for (int k = 0; k < dim; k++) last -= l[k];
    l[dim] = last;
    const double w = data[q][3];
    m0 += w;
    for (int a = 0; a < nv; a++) {
      m1[a] += w * l[a];
      for (int b = 0; b < nv; b++) {
        m2[a][b] += w * l[a] * l[b];
        for (int c = 0; c < nv; c++) {
          m3[a][b][c] += w * l[a] * l[b] * l[c];
          for (int d = 0; d < nv; d++) m4[a][b][c][d] += w * l[a] * l[b] * l[c] * l[d];
        }
      }
    }
  }
  RuleMom M{};
  M.m0 = m0; M.m1 = m1[0]; M.m2s = m2[0][0]; M.m2d = m2[0][1]; M.m3aab = m3[0][0][1]; M.m3abc = m3[0][1][2];
  M.m4aabb = m4[0][0][1][1]; M.m4aabc = m4[0][0][1][2]; M.m4abcd = dim == 3 ? m4[0][1][2][3] : 0.0;
  const double tol = 1e-13 * fabs(m0);
  bool sym = true;
  auto chk = [&](double v, double ref) { if (fabs(v - ref) > tol) sym = false; };
  for (int a = 0; a < nv; a++) {
    chk(m1[a], M.m1);
    for (int b = 0; b < nv; b++) {
      chk(m2[a][b], a == b ? M.m2s : M.m2d);
      for (int c = 0; c < nv; c++) {
        const int distinct3 = (a != b) + (a != c) + (b != c);          // 0: aaa, 2: aab, 3: abc
        if (distinct3 == 2) chk(m3[a][b][c], M.m3aab);
        if (distinct3 == 3) chk(m3[a][b][c], M.m3abc);
        for (int d = 0; d < nv; d++) {
          int v[4] = {a, b, c, d};
          std::sort(v, v + 4);
          const bool aabb = v[0] == v[1] && v[2] == v[3] && v[1] != v[2];
          const bool allx = v[0] != v[1] && v[1] != v[2] && v[2] != v[3];
          const bool aabc = !aabb && !allx && ((v[0] == v[1]) + (v[1] == v[2]) + (v[2] == v[3]) == 1);
          if (aabb) chk(m4[a][b][c][d], M.m4aabb);
          if (aabc) chk(m4[a][b][c][d], M.m4aabc);
          if (allx) chk(m4[a][b][c][d], M.m4abcd);
        }
      }
    }
  }
  *out = M;
  return sym;
}

// ---- the row kernel -------------------------------------------------------------------------------------------------
struct RowArgs {
  int64_t ntiles;
  const TileHdr *tiles;
  const uint4 *desc;
  const double *G;
  const double *cutmats;
  double *vals;
  unsigned *counter;       // tile claim counter, zeroed before the launch
  int accumulate;
  // rule constants of the closed forms (row_vertex / row_edge)
  double m0, m2s, m2d, hm1, q3aab, q3abc, qm2d, qdm, q4[3];
  double q3d, q4a, q4b;    // mask forms: q3aab - q3abc; q4(cnt) = q4[0] + cnt (q4a + q4b cnt), cnt = shared vertices of two edges
};

template <int J> __device__ __forceinline__ int desc_pos(const uint4 &d) {
  constexpr int b = 6 + J, w = b >> 2, sh = (b & 3) * 8;
  const uint32_t x = w == 1 ? d.y : (w == 2 ? d.z : d.w);
  return (int)((x >> sh) & 0xffu);
}
template <int ND, int J> __device__ __forceinline__ void desc_positions(const uint4 &d, int *pos) {
  if constexpr (J < ND) { pos[J] = desc_pos<J>(d); desc_positions<ND, J + 1>(d, pos); }
}

// 32-byte global load (sm_100: LDG.E.256)
__device__ __forceinline__ void ldg256(const double *p, double *o) {
  asm volatile("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(o[0]), "=d"(o[1]), "=d"(o[2]), "=d"(o[3]) : "l"(p));
}

// Row i of the constant-rule element matrix  A_ij = sum_r w_r (grad phi_i)^T Ghat (grad phi_j) + gm sum_r w_r phi_i phi_j,
// Ghat = |J| c_s J^-1 J^-T, gm = |J| c_m, for a permutation-symmetric rule.  With K[a][b] = grad(lam_a) Ghat grad(lam_b):
//   vertex i, vertex c:       m0 K[i][c] + gm M2(i,c)                         M2 = m2s (same vertex) / m2d
//   vertex i, edge (c,d):    -1/2 m1 (K[i][c] + K[i][d]) - 1/2 gm M3(i,c,d)   M3 = m3aab (i on the edge) / m3abc
//   edge (a,b), vertex c:    -1/2 m1 (K[a][c] + K[b][c]) - 1/2 gm M3(c,a,b)
//   edge (a,b), edge (c,d):   1/4 [K[b][d] M2(a,c) + K[b][c] M2(a,d) + K[a][d] M2(b,c) + K[a][c] M2(b,d)] + 1/4 gm M4(a,b,c,d)
// (phi_edge = -1/2 lam_a lam_b, NGS/fem/recursive_pol.hpp:1372).  ka = K[a][0..DIM-1] is what k_geom stores per local vertex;
// K[a][DIM] = -sum.  Same quadrature as the reference's static rule, re-associated.
template <int DIM, int ORDER>
__device__ __forceinline__ void row_vertex(const RowArgs &A, int i, const double *ga, double *v) {
  constexpr int NEDGE = H1<DIM, ORDER>::NEDGE;
  double kv[DIM + 1];
  double s = 0.0;
#pragma unroll
  for (int c = 0; c < DIM; c++) { kv[c] = ga[c]; s += ga[c]; }
  kv[DIM] = -s;
  const double gm = ga[3];
  const double gs = gm * A.m2s, gd = gm * A.m2d;
#pragma unroll
  for (int c = 0; c <= DIM; c++) v[c] = fma(A.m0, kv[c], i == c ? gs : gd);
  if constexpr (ORDER == 2) {
    const double gin = gm * A.q3aab, gout = gm * A.q3abc;
#pragma unroll
    for (int e = 0; e < NEDGE; e++) {
      int c, d;
      edge_vertices<DIM>(e, c, d);
      v[DIM + 1 + e] = fma(A.hm1, kv[c] + kv[d], (i == c || i == d) ? gin : gout);
    }
  }
}

// vertices of edge e from packed 2-bit tables (edge_vertices<DIM> for a run-time e without a select chain)
template <int DIM> __device__ __forceinline__ void edge_vertices_rt(int e, int &a, int &b) {
  constexpr uint32_t TA = DIM == 2 ? 0x6u : 0x43fu, TB = DIM == 2 ? 0x18u : 0xa64u;
  a = (int)((TA >> (2 * e)) & 3u);
  b = (int)((TB >> (2 * e)) & 3u);
}

// Edge row (a, b).  a and b are run-time values (the local row of the pair), the columns are compile-time: instead of
// selecting K entries by comparing indices (two FSEL per double select, ~55 of them), the Kronecker deltas are 0.0 / 1.0
// masks fa[c] = [c == a], fb[c] = [c == b] and the selects become fused multiply-adds on the otherwise idle FP64 pipe:
//   corr(c, d) = fa[c] kb[d] + fa[d] kb[c] + fb[c] ka[d] + fb[d] ka[c]         (the M2(x, y) = m2d + (m2s - m2d) [x == y] part)
//   M3, M4 terms: polynomials in fin[c] = fa[c] + fb[c] (0 / 1) and cnt = fin[c] + fin[d] (0 / 1 / 2), exact at those values
//   up to one rounding of the host-computed coefficients.
template <int DIM>
__device__ __forceinline__ void row_edge(const RowArgs &A, int a, int b, const double *ga, const double *gb, double *v) {
  constexpr int NEDGE = H1<DIM, 2>::NEDGE;
  double ka[DIM + 1], kb[DIM + 1], ks[DIM + 1], fa[DIM + 1], fb[DIM + 1], fin[DIM + 1];
  double sa = 0.0, sb = 0.0;
#pragma unroll
  for (int c = 0; c < DIM; c++) { ka[c] = ga[c]; kb[c] = gb[c]; sa += ga[c]; sb += gb[c]; }
  ka[DIM] = -sa; kb[DIM] = -sb;
#pragma unroll
  for (int c = 0; c <= DIM; c++) {
    ks[c] = ka[c] + kb[c];
    fa[c] = __hiloint2double(c == a ? 0x3ff00000 : 0, 0);
    fb[c] = __hiloint2double(c == b ? 0x3ff00000 : 0, 0);
    fin[c] = fa[c] + fb[c];
  }
  const double gm = ga[3];
  const double gd3 = gm * A.q3d, gout = gm * A.q3abc;
#pragma unroll
  for (int c = 0; c <= DIM; c++) v[c] = fma(A.hm1, ks[c], fma(fin[c], gd3, gout));
  const double g0 = gm * A.q4[0], g4a = gm * A.q4a, g4b = gm * A.q4b;
#pragma unroll
  for (int e = 0; e < NEDGE; e++) {
    int c, d;
    edge_vertices<DIM>(e, c, d);
    const double corr = fma(fa[c], kb[d], fma(fa[d], kb[c], fma(fb[c], ka[d], fb[d] * ka[c])));
    const double cnt = fin[c] + fin[d];
    const double gq = fma(cnt, fma(cnt, g4b, g4a), g0);
    v[DIM + 1 + e] = fma(A.qm2d, ks[c] + ks[d], fma(A.qdm, corr, gq));
  }
}

// row i of a quadrature (cut) element matrix: nd contiguous doubles of the cut buffer
template <int ND>
__device__ __forceinline__ void row_values_cut(const double *cm, double *v) {
  if constexpr (ND % 2 == 0) {
#pragma unroll
    for (int j = 0; j < ND; j += 2) {
      const double2 t = __ldg(reinterpret_cast<const double2 *>(cm + j));
      v[j] = t.x; v[j + 1] = t.y;
    }
  } else {
#pragma unroll
    for (int j = 0; j < ND; j++) v[j] = __ldg(cm + j);
  }
}

// x * x the optimiser cannot see through (it knows bit 1 of a square is clear)
__device__ __forceinline__ uint32_t opaque_square(uint32_t x) {
  uint32_t y;
  asm volatile("mul.lo.u32 %0, %1, %1;" : "=r"(y) : "r"(x));
  return y;
}

// geometry record(s) of a pair: vertex row i -> record i; edge row (a, b) -> records a and b (the same 128-byte line).
// Returns 0 -- computed from the high bits of the record address, which the compiler cannot fold: added to the index of the
// descriptor load that follows, it keeps that load BEHIND the address arithmetic here.  (Hoisted above it, the arithmetic
// reuses registers of earlier loads and its scoreboard wait then also waits for the descriptor load just issued.)
template <int DIM, int ORDER>
__device__ __forceinline__ uint32_t load_geo(const RowArgs &A, uint32_t w0, double *ga, double *gb) {
  constexpr int GS = (DIM + 1) * 4;
  const int i = (int)(w0 >> DESC_KBITS);
  const double *Gk = A.G + (int64_t)(w0 & DESC_KMASK) * GS;
  int a = i, b = i;
  if (ORDER == 2 && i > DIM) edge_vertices_rt<DIM>(i - (DIM + 1), a, b);
  const double *pa = Gk + 4 * a;
  ldg256(pa, ga);
  if (ORDER == 2 && i > DIM) ldg256(Gk + 4 * b, gb);
  return (uint32_t)(reinterpret_cast<uintptr_t>(pa) >> 60);
}

// Row of a quadrature (cut) element: nd doubles of the cut buffer added to the window.  Out of line on purpose -- inlined, its
// loads take the scoreboard of the next step's descriptor / geometry loads and every step of the fast path waits for those.
template <int ND, bool LOWER>
__device__ __noinline__ void cut_row_add(const double *__restrict__ cm, double *accrow, uint4 d) {
  double v[ND];
  int pos[ND];
  row_values_cut<ND>(cm, v);
  desc_positions<ND, 0>(d, pos);
#pragma unroll
  for (int j = 0; j < ND; j++)
    if (!LOWER || pos[j] != POS_SKIP) accrow[pos[j]] += v[j];
}

// one (row, element) pair with a constant rule: closed-form row added at the pair's column positions of the window
template <int DIM, int ORDER, bool LOWER>
__device__ __forceinline__ void row_step(const RowArgs &A, double *accrow, const uint4 &d, const double *ga, const double *gb) {
  constexpr int ND = H1<DIM, ORDER>::ND;
  const int i = (int)(d.x >> DESC_KBITS);
  double v[ND];
  if (ORDER == 1 || i <= DIM) {
    row_vertex<DIM, ORDER>(A, i, ga, v);
  } else {
    if constexpr (ORDER == 2) {
      int ea, eb;
      edge_vertices_rt<DIM>(i - (DIM + 1), ea, eb);
      row_edge<DIM>(A, ea, eb, ga, gb, v);
    }
  }
  int pos[ND];
  desc_positions<ND, 0>(d, pos);
  // the nd columns of an element are distinct: read a batch, add, write it (no serialised read-modify-write chain; two
  // batches keep the live registers under the 128 of four CTAs per SM)
  if constexpr (LOWER) {                                          // symmetric storage: columns above the row have no slot
#pragma unroll
    for (int j = 0; j < ND; j++)
      if (pos[j] != POS_SKIP) accrow[pos[j]] += v[j];
    return;
  }
  constexpr int HB = (ND + 1) / 2;
  double w[HB];
#pragma unroll
  for (int j = 0; j < HB; j++) w[j] = accrow[pos[j]];
#pragma unroll
  for (int j = 0; j < HB; j++) accrow[pos[j]] = w[j] + v[j];
#pragma unroll
  for (int j = HB; j < ND; j++) w[j - HB] = accrow[pos[j]];
#pragma unroll
  for (int j = HB; j < ND; j++) accrow[pos[j]] = w[j - HB] + v[j];
}

// 64-byte tile header global -> the warp's shared slot, asynchronously (lanes 0..3 move 16 bytes each)
__device__ __forceinline__ void header_fetch(const TileHdr *src, TileHdr *slot, int lane) {
  if (lane < 4) {
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(reinterpret_cast<char *>(slot) + 16 * lane);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(reinterpret_cast<const char *>(src) + 16 * lane));
  }
  asm volatile("cp.async.commit_group;");
}
__device__ __forceinline__ void header_wait() {
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncwarp();
}

// WARP = one tile of consecutive CSR rows whose values (<= ACCW doubles) live in the warp's private shared-memory window:
// initialised by the warp (zero, or the current values when accumulating), gathered, and streamed out once with coalesced
// stores.  Warps never wait for each other; they claim tiles from a counter in plan order (neighbouring tiles run at about
// the same time: shared geometry records are L2 hits; no tail), one tile ahead so that the next header is in shared memory
// and the first descriptors of the next tile fly while the window of this one is written out.
// LANE = the rows the schedule dealt to it: per step one (row, element) pair.
// Software pipeline in registers: at the top of step s the geometry of step s and the descriptors of steps s, s+1, s+2 have
// landed; the geometry of step s+1 and the descriptor of step s+3 are issued there, and nothing else in the step reads or
// writes their registers.  All global loads share one hardware scoreboard (a counter), so ANY wait on it after that point --
// a register copy, an address computed into a register an earlier load wrote -- waits for the loads just issued as well.
// Hence: the register rotation and the first read of this step's geometry sit at the very top (kept there by a data
// dependence the optimiser cannot fold), the descriptor load is pinned behind the geometry address arithmetic, the
// quadrature rows are out of line and last in the step (prefetched into L1 at the top), and two steps per trip let the two
// geometry sets alternate without copies.
// No two lanes share a row and the nd columns of one element are distinct, so every CSR entry is summed in plan order.
#define NGSX_ROW_STEP(S, GA, GB, NA, NB)                                                   \
  {                                                                                        \
    const uint32_t ghi = (uint32_t)__double2hiint(GA[3]);                                  \
    uint32_t zero = opaque_square(ghi | d2.x) & 2u;  /* squares are 0 or 1 mod 4 */        \
    const uint4 d = d0;                                                                    \
    d0 = d1; d1 = d2;                                                                      \
    if ((S) + 1 < myload) zero = load_geo<DIM, ORDER>(A, d0.x + zero, NA, NB);             \
    ldg128_if((S) + 3 < myload, A.desc + base3 + lane + zero, d2);                         \
    base3 += __popc(__ballot_sync(0xffffffffu, (S) + 3 < myload));                         \
    if ((S) < myload) {                                                                    \
      double *accrow = acc + (d.y & 0xffffu);                                              \
      const bool coded = GA[3] != GA[3];                                /* NaN-boxed code */ \
      const unsigned payload = (unsigned)((unsigned long long)__double_as_longlong(GA[3]) & 0xffffffffULL); \
      const bool cut = coded && payload != 0u;                                             \
      const double *cm = A.cutmats + (int64_t)(payload - 1u) * NE + (int)(d.x >> DESC_KBITS) * ND; \
      if (cut) { prefetch_l1(cm); prefetch_l1(cm + ND - 1); }         /* lands during the closed-form rows of the other lanes */ \
      if (!coded) row_step<DIM, ORDER, LOWER>(A, accrow, d, GA, GB);                       \
      if (cut) cut_row_add<ND, LOWER>(cm, accrow, d);                                      \
    }                                                                                      \
  }

// descriptors of steps 0, 1, 2 of a tile
#define NGSX_ROW_PRIME()                                                                   \
  {                                                                                        \
    d0 = null4; d1 = null4; d2 = null4;                                                    \
    ldg128_if(0 < myload, A.desc + base3 + lane, d0);                                      \
    base3 += __popc(__ballot_sync(0xffffffffu, 0 < myload));                               \
    ldg128_if(1 < myload, A.desc + base3 + lane, d1);                                      \
    base3 += __popc(__ballot_sync(0xffffffffu, 1 < myload));                               \
    ldg128_if(2 < myload, A.desc + base3 + lane, d2);                                      \
    base3 += __popc(__ballot_sync(0xffffffffu, 2 < myload));                               \
  }

template <int DIM, int ORDER, bool LOWER>
__global__ void __launch_bounds__(RW * 32, NGSX_MINBLOCKS) k_rows(RowArgs A) {
  constexpr int ND = H1<DIM, ORDER>::ND, NE = ND * ND;
  extern __shared__ __align__(16) double smem_rows[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *acc = smem_rows + warp * ACCW;
  TileHdr *slot = reinterpret_cast<TileHdr *>(smem_rows + RW * ACCW) + warp;
  const uint4 null4 = make_uint4(0xffffffffu, 0u, 0u, 0u);
  unsigned tt = 0;
  if (lane == 0) tt = atomicAdd(A.counter, 1u);
  tt = __shfl_sync(0xffffffffu, tt, 0);
  if (tt >= A.ntiles) return;
  header_fetch(A.tiles + tt, slot, lane);
  header_wait();
  int64_t z0 = slot->z0, base3 = slot->d0;
  int zlen = slot->zlen, nsteps = slot->nsteps, myload = slot->load[lane];
  uint4 d0, d1, d2;
  NGSX_ROW_PRIME()
  for (;;) {
    __syncwarp();                                                  // every lane has read the slot
    unsigned nxt = 0;
    if (lane == 0) nxt = atomicAdd(A.counter, 1u);
    nxt = __shfl_sync(0xffffffffu, nxt, 0);
    const bool more = nxt < A.ntiles;
    if (more) header_fetch(A.tiles + nxt, slot, lane);
    if (A.accumulate) for (int t = lane; t < zlen; t += 32) acc[t] = A.vals[z0 + t];
    else for (int t = lane; t < zlen; t += 32) acc[t] = 0.0;
    double ga[4] = {0.0, 0.0, 0.0, 0.0}, gb[4] = {0.0, 0.0, 0.0, 0.0}, ha[4] = {0.0, 0.0, 0.0, 0.0}, hb[4] = {0.0, 0.0, 0.0, 0.0};
    if (0 < myload) load_geo<DIM, ORDER>(A, d0.x, ga, gb);
    __syncwarp();
#if NGSX_ROWS_UNROLL == 2
    for (int s = 0; s < nsteps; s += 2) {
      NGSX_ROW_STEP(s, ga, gb, ha, hb)
      if (s + 1 >= nsteps) break;
      NGSX_ROW_STEP(s + 1, ha, hb, ga, gb)
    }
#else
    for (int s = 0; s < nsteps; s++) {
      NGSX_ROW_STEP(s, ga, gb, ha, hb)
#pragma unroll
      for (int m = 0; m < 4; m++) { ga[m] = ha[m]; gb[m] = hb[m]; }
    }
#endif
    __syncwarp();
    const int64_t zc = z0;
    const int zl = zlen;
    if (more) {                                                    // the next tile's first descriptors fly during the write-out
      header_wait();
      z0 = slot->z0; base3 = slot->d0;
      zlen = slot->zlen; nsteps = slot->nsteps; myload = slot->load[lane];
      NGSX_ROW_PRIME()
    }
    for (int t = lane; t < zl; t += 32) A.vals[zc + t] = acc[t];
    if (!more) break;
  }
}
#undef NGSX_ROW_STEP
#undef NGSX_ROW_PRIME

template <int DIM, int ORDER, bool LOWER = false>
static int launch_rows_t(const RowArgs &A, int sm_count, cudaStream_t st) {
  size_t smem = (size_t)RW * (ACCW * sizeof(double) + sizeof(TileHdr));
  auto kern = k_rows<DIM, ORDER, LOWER>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, RW * 32, smem));
  if (occ < 1) { set_error("row kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
  int64_t need = (A.ntiles + RW - 1) / RW;
  unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)sm_count * occ));
  NGSX_CUDA(cudaMemsetAsync(A.counter, 0, sizeof(unsigned), st));
  kern<<<grid, RW * 32, smem, st>>>(A);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int rows_reserve_cut(ngsx_rowplan *P, int64_t cap) {
  cap = std::max<int64_t>(1, std::min<int64_t>(cap, std::max<int64_t>(1, P->nlist)));
  if (cap <= P->cut_cap) return NGSX_OK;
  if (P->cutmats) { NGSX_CUDA(cudaDeviceSynchronize()); cudaFree(P->cutmats); P->cutmats = nullptr; P->cut_cap = 0; }
  NGSX_CUDA(cudaMalloc(&P->cutmats, (size_t)cap * P->nd * P->nd * sizeof(double)));
  P->cut_cap = cap;
  return NGSX_OK;
}

int rows_scratch(ngsx_rowplan *P, int dim) {
  if (P->G && P->g_dim != dim) { NGSX_CUDA(cudaDeviceSynchronize()); cudaFree(P->G); P->G = nullptr; }
  if (!P->G) { NGSX_CUDA(cudaMalloc(&P->G, std::max<int64_t>(1, P->nlist) * (dim + 1) * 4 * sizeof(double))); P->g_dim = dim; }
  if (!P->cutlist) NGSX_CUDA(cudaMalloc(&P->cutlist, std::max<int64_t>(1, P->nlist) * sizeof(int32_t)));
  if (!P->cutcount) NGSX_CUDA(cudaMalloc(&P->cutcount, sizeof(int32_t)));
  if (!P->counter) NGSX_CUDA(cudaMalloc(&P->counter, sizeof(unsigned)));
  return NGSX_OK;
}

int launch_rows(const ngsx_rowplan *P, int dim, const RuleMom &M, int accumulate, double *vals, int sm_count, cudaStream_t st) {
  if (P->nrows == 0) return NGSX_OK;
  if (P->ds != 16) { set_error("row gather: this plan was built for merged facet-patch dof lists (use ngsx_assemble_csr_rows_mats)"); return NGSX_ERR_ARG; }
  RowArgs A{};
  A.ntiles = P->ntiles; A.tiles = P->tiles; A.desc = reinterpret_cast<const uint4 *>(P->desc); A.G = P->G; A.cutmats = P->cutmats;
  A.vals = vals; A.accumulate = accumulate; A.counter = P->counter;
  A.m0 = M.m0; A.m2s = M.m2s; A.m2d = M.m2d; A.hm1 = -0.5 * M.m1; A.q3aab = -0.5 * M.m3aab; A.q3abc = -0.5 * M.m3abc;
  A.qm2d = 0.25 * M.m2d; A.qdm = 0.25 * (M.m2s - M.m2d);
  A.q4[0] = 0.25 * M.m4abcd; A.q4[1] = 0.25 * M.m4aabc; A.q4[2] = 0.25 * M.m4aabb;
  A.q3d = A.q3aab - A.q3abc;
  A.q4b = 0.5 * (A.q4[2] - 2.0 * A.q4[1] + A.q4[0]);
  A.q4a = A.q4[1] - A.q4[0] - A.q4b;
  if (P->lower)
    switch (dim * 100 + P->nd) {
      case 203: return launch_rows_t<2, 1, true>(A, sm_count, st);
      case 206: return launch_rows_t<2, 2, true>(A, sm_count, st);
      case 304: return launch_rows_t<3, 1, true>(A, sm_count, st);
      case 310: return launch_rows_t<3, 2, true>(A, sm_count, st);
    }
  switch (dim * 100 + P->nd) {
    case 203: return launch_rows_t<2, 1>(A, sm_count, st);
    case 206: return launch_rows_t<2, 2>(A, sm_count, st);
    case 304: return launch_rows_t<3, 1>(A, sm_count, st);
    case 310: return launch_rows_t<3, 2>(A, sm_count, st);
  }
  set_error("row gather: unsupported (dofs per element, dimension)");
  return NGSX_ERR_UNSUPPORTED;
}

}  // namespace ngsx

using namespace ngsx;

extern "C" {

int ngsx_rowplan_destroy(ngsx_rowplan *P) {
  if (!P) return NGSX_OK;
  cudaDeviceSynchronize();
  cudaFree(P->tiles); cudaFree(P->desc); cudaFree(P->rowslot);
  cudaFree(P->G); cudaFree(P->cutlist); cudaFree(P->cutcount); cudaFree(P->cutmats); cudaFree(P->counter);
  delete P;
  return NGSX_OK;
}

int ngsx_rowplan_info(const ngsx_rowplan *P, int64_t *h_info) {
  NGSX_REQUIRE(P && h_info, "rowplan_info: bad arguments");
  h_info[0] = P->nrows; h_info[1] = P->npairs; h_info[2] = P->nnz_covered; h_info[3] = P->max_row;
  h_info[4] = P->nnz_lo; h_info[5] = P->nnz_hi;
  return NGSX_OK;
}

int ngsx_rowplan_cut_usage(ngsx_rowplan *P, int64_t *h_count, int64_t *h_cap, void *stream) {
  NGSX_REQUIRE(P && h_count && h_cap, "rowplan_cut_usage: bad arguments");
  *h_count = 0; *h_cap = P->cut_cap;
  if (!P->cutcount) return NGSX_OK;
  int32_t c = 0;
  NGSX_CUDA(cudaMemcpyAsync(&c, P->cutcount, sizeof c, cudaMemcpyDeviceToHost, as_stream(stream)));
  NGSX_CUDA(cudaStreamSynchronize(as_stream(stream)));
  *h_count = c;
  return NGSX_OK;
}

int ngsx_rowplan_reserve_cut(ngsx_rowplan *P, int64_t capacity) {
  NGSX_REQUIRE(P && capacity >= 0, "rowplan_reserve_cut: bad arguments");
  return rows_reserve_cut(P, capacity);
}

int ngsx_rowplan_rowslots(ngsx_rowplan *P, const int32_t **d_rowslot, int64_t *h_nslots, void *stream) {
  NGSX_REQUIRE(P && d_rowslot && h_nslots, "rowplan_rowslots: bad arguments");
  cudaStream_t st = as_stream(stream);
  if (!P->rowslot) {
    const int64_t n = std::max<int64_t>(1, P->nlist * P->nd);
    NGSX_CUDA(cudaMalloc(&P->rowslot, n * sizeof(int32_t)));
    NGSX_CUDA(cudaMemsetAsync(P->rowslot, 0xff, n * sizeof(int32_t), st));
    if (P->nslots > 0) {
      k_rowslots<<<(unsigned)((P->nslots + 255) / 256), 256, 0, st>>>(P->nslots, P->nd, P->ds, P->desc, P->rowslot);
      NGSX_LAUNCH_CHECK();
    }
    NGSX_CUDA(cudaStreamSynchronize(st));
  }
  *d_rowslot = P->rowslot; *h_nslots = P->nslots;
  return NGSX_OK;
}

static int rowplan_create(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                          int64_t row_lo, int64_t row_hi, const int32_t *row_owner, int32_t me, ngsx_rowplan **out, void *stream, int lower);

int ngsx_rowplan_create(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                        int64_t row_lo, int64_t row_hi, ngsx_rowplan **out, void *stream) {
  return rowplan_create(nlist, nd, table, ndof, rowptr, colidx, row_lo, row_hi, nullptr, 0, out, stream, 0);
}

// the same plan over a LOWER-TRIANGLE pattern (symmetric storage, AddElementMatrixSymmetric, NGS/linalg/sparsematrix_impl.hpp:605-690):
// rowptr / colidx hold the columns <= row only; the local columns of a pair that lie above its row are skipped by the gather
int ngsx_rowplan_create_lower(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                              int64_t row_lo, int64_t row_hi, ngsx_rowplan **out, void *stream) {
  return rowplan_create(nlist, nd, table, ndof, rowptr, colidx, row_lo, row_hi, nullptr, 0, out, stream, 1);
}

int ngsx_rowplan_create_owned(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                              const int32_t *row_owner, int32_t me, ngsx_rowplan **out, void *stream) {
  if (!row_owner) { set_error("rowplan_create_owned: row_owner is NULL"); return NGSX_ERR_ARG; }
  return rowplan_create(nlist, nd, table, ndof, rowptr, colidx, 0, -1, row_owner, me, out, stream, 0);
}

static int rowplan_create(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                          int64_t row_lo, int64_t row_hi, const int32_t *row_owner, int32_t me, ngsx_rowplan **out, void *stream, int lower) {
  NGSX_REQUIRE(out && nlist >= 0 && nd > 0 && nd <= 26 && ndof >= 0 && ndof < (1LL << 31) && rowptr, "rowplan_create: bad arguments");
  NGSX_REQUIRE(nlist == 0 || (table && colidx), "rowplan_create: table / colidx are NULL");
  NGSX_REQUIRE(nlist * nd < (1LL << 31), "rowplan_create: more than 2^31 (element, dof) pairs");
  NGSX_REQUIRE(nlist <= (int64_t)DESC_KMASK, "rowplan_create: more than 2^27 - 1 list entries in one plan");
  cudaStream_t st = as_stream(stream);
  if (row_hi < 0 || row_hi > ndof) row_hi = ndof;
  if (row_lo < 0) row_lo = 0;
  ngsx_rowplan *P = new ngsx_rowplan();
  P->nlist = nlist; P->nd = nd; P->ndof = ndof; P->ds = nd <= 10 ? 16 : 32; P->rowptr = rowptr; P->lower = lower;
  *out = P;
  const int64_t nall = nlist * nd;
  if (nall == 0 || row_hi <= row_lo) return NGSX_OK;

  uint64_t *keys = nullptr, *alt = nullptr;
  uint8_t *flags = nullptr;
  void *tmp = nullptr;
  int64_t *d_scalar = nullptr;          // [0] valid pairs, [1] rows
  unsigned long long *d_stat = nullptr;
  int *d_bad = nullptr;
  int64_t *d_pairptr = nullptr, *d_span = nullptr, *d_tile_d0 = nullptr, *d_tile_so = nullptr;
  RowSched *d_sched = nullptr;
  int32_t *d_stepoffs = nullptr, *d_order = nullptr, *d_ids = nullptr;
  uint32_t *d_ck = nullptr, *d_ck2 = nullptr;
  TileHdr *d_tiles_nat = nullptr;
  auto cleanup = [&]() {
    cudaFree(keys); cudaFree(alt); cudaFree(flags); cudaFree(tmp); cudaFree(d_scalar); cudaFree(d_stat); cudaFree(d_bad);
    cudaFree(d_pairptr); cudaFree(d_span); cudaFree(d_tile_d0); cudaFree(d_tile_so); cudaFree(d_sched); cudaFree(d_stepoffs);
    cudaFree(d_order); cudaFree(d_ids); cudaFree(d_ck); cudaFree(d_ck2); cudaFree(d_tiles_nat);
  };
#define PLAN_CUDA(call)                                                                                         \
  do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { set_error(std::string(#call) + ": " + cudaGetErrorString(e__)); \
       cleanup(); ngsx_rowplan_destroy(P); *out = nullptr; return NGSX_ERR_CUDA; } } while (0)
#define PLAN_FAIL(status, msg) do { set_error(msg); cleanup(); ngsx_rowplan_destroy(P); *out = nullptr; return status; } while (0)

  PLAN_CUDA(cudaMalloc(&keys, nall * sizeof(uint64_t)));
  PLAN_CUDA(cudaMalloc(&alt, nall * sizeof(uint64_t)));
  PLAN_CUDA(cudaMalloc(&d_scalar, 2 * sizeof(int64_t)));
  PLAN_CUDA(cudaMalloc(&d_stat, 4 * sizeof(unsigned long long)));
  PLAN_CUDA(cudaMalloc(&d_bad, sizeof(int)));
  PLAN_CUDA(cudaMemsetAsync(d_stat, 0, 4 * sizeof(unsigned long long), st));
  PLAN_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
  k_pair_keys<<<(unsigned)((nall + 255) / 256), 256, 0, st>>>(nall, nd, nlist, table, row_lo, row_hi, row_owner, me, keys);
  cub::DoubleBuffer<uint64_t> db(keys, alt);
  size_t tmp_bytes = 0;
  PLAN_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, db, nall, 0, 64, st));
  PLAN_CUDA(cudaMalloc(&tmp, tmp_bytes));
  PLAN_CUDA(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, db, nall, 0, 64, st));
  const uint64_t *sorted = db.Current();
  k_lower_bound<<<1, 1, 0, st>>>(nall, sorted, (uint64_t)row_hi << 32, d_scalar);
  int64_t npairs = 0;
  PLAN_CUDA(cudaMemcpyAsync(&npairs, d_scalar, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaStreamSynchronize(st));
  g_launches.fetch_add(4);
  P->npairs = npairs; P->nslots = npairs;
  if (npairs == 0) { cleanup(); return NGSX_OK; }

  PLAN_CUDA(cudaMalloc(&flags, npairs));
  k_head_flags<<<(unsigned)((npairs + 255) / 256), 256, 0, st>>>(npairs, sorted, flags);
  const int64_t cap_rows = std::min<int64_t>(npairs, row_hi - row_lo);
  PLAN_CUDA(cudaMalloc(&d_pairptr, (cap_rows + 1) * sizeof(int64_t)));
  cudaFree(tmp); tmp = nullptr; tmp_bytes = 0;
  thrust::counting_iterator<int64_t> counting(0);
  PLAN_CUDA(cub::DeviceSelect::Flagged(nullptr, tmp_bytes, counting, flags, d_pairptr, d_scalar + 1, npairs, st));
  PLAN_CUDA(cudaMalloc(&tmp, tmp_bytes));
  PLAN_CUDA(cub::DeviceSelect::Flagged(tmp, tmp_bytes, counting, flags, d_pairptr, d_scalar + 1, npairs, st));
  int64_t nrows = 0;
  PLAN_CUDA(cudaMemcpyAsync(&nrows, d_scalar + 1, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaStreamSynchronize(st));
  g_launches.fetch_add(3);
  P->nrows = nrows;
  PLAN_CUDA(cudaMalloc(&d_span, 2 * nrows * sizeof(int64_t)));
  k_plan_rows<<<(unsigned)((nrows + 255) / 256), 256, 0, st>>>(nrows, npairs, sorted, d_pairptr, rowptr, d_span, d_stat);
  g_launches.fetch_add(1);
  std::vector<int64_t> h_span(2 * nrows), h_pairptr(nrows + 1);
  unsigned long long stat[4];
  PLAN_CUDA(cudaMemcpyAsync(h_span.data(), d_span, 2 * nrows * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaMemcpyAsync(h_pairptr.data(), d_pairptr, (nrows + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaMemcpyAsync(stat, d_stat, sizeof stat, cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaStreamSynchronize(st));
  PLAN_CUDA(cudaGetLastError());
  P->nnz_covered = (int64_t)stat[0]; P->max_row = (int64_t)stat[1]; P->nnz_lo = (int64_t)stat[2]; P->nnz_hi = (int64_t)stat[3];
  if (P->max_row > MAXROW)
    PLAN_FAIL(NGSX_ERR_UNSUPPORTED, "rowplan_create: a CSR row is longer than 256 entries; use the coloured scatter (ngsx_assemble_csr_h1)");

  // ---- the schedule, on the host (symbolic phase: a sequential pass over the plan rows) ----
  // Tiles: greedy over the plan rows; a tile is closed when its CSR span would exceed the warp's window, at TILE_MAXROWS rows
  // or at TILE_MAXPAIRS pairs.  Lanes: rows by decreasing pair count to the least loaded lane (ties: the lower lane), then
  // the lanes are renumbered by decreasing load so that the active lanes of every step form a prefix.
  std::vector<TileHdr> h_tiles;
  std::vector<int64_t> h_tile_so;                 // first entry of the tile in h_stepoffs
  std::vector<int32_t> h_stepoffs;                // per tile and step: descriptors before that step
  std::vector<RowSched> h_sched(nrows);
  h_tiles.reserve(nrows / 16 + 2);
  int64_t dcount = 0, max_steps = 0;
  bool too_many = false;
  {
    std::vector<int> order, lane_of(TILE_MAXROWS), start_of(TILE_MAXROWS), rank(32);
    int64_t q0 = 0;
    auto close_tile = [&](int64_t qa, int64_t qb) {
      const int n = (int)(qb - qa);
      order.resize(n);
      std::iota(order.begin(), order.end(), 0);
      std::stable_sort(order.begin(), order.end(), [&](int x, int y) {
        return h_pairptr[qa + x + 1] - h_pairptr[qa + x] > h_pairptr[qa + y + 1] - h_pairptr[qa + y]; });
      int load[32] = {0};
      for (int t = 0; t < n; t++) {
        const int r = order[t];
        int best = 0;
        for (int l = 1; l < 32; l++) if (load[l] < load[best]) best = l;
        lane_of[r] = best; start_of[r] = load[best];
        load[best] += (int)(h_pairptr[qa + r + 1] - h_pairptr[qa + r]);
      }
      int perm[32];
      std::iota(perm, perm + 32, 0);
      std::stable_sort(perm, perm + 32, [&](int x, int y) { return load[x] > load[y]; });
      for (int l = 0; l < 32; l++) rank[perm[l]] = l;
      TileHdr H{};
      H.z0 = h_span[2 * qa];
      H.zlen = (int32_t)(h_span[2 * (qb - 1) + 1] - H.z0);
      H.d0 = dcount;
      H.nsteps = load[perm[0]];
      if (H.nsteps > 255) too_many = true;
      for (int l = 0; l < 32; l++) H.load[l] = (uint8_t)std::min(255, load[perm[l]]);
      const int32_t tile = (int32_t)h_tiles.size();
      h_tile_so.push_back((int64_t)h_stepoffs.size());
      int32_t acc = 0;
      for (int s = 0; s < H.nsteps; s++) {
        h_stepoffs.push_back(acc);
        int cnt = 0;
        for (int l = 0; l < 32; l++) cnt += load[l] > s;
        acc += cnt;
      }
      dcount += acc;
      max_steps = std::max<int64_t>(max_steps, H.nsteps);
      for (int r = 0; r < n; r++) {
        RowSched rs;
        rs.tile = tile; rs.start = start_of[r]; rs.lane = rank[lane_of[r]]; rs.off = (int32_t)(h_span[2 * (qa + r)] - H.z0);
        h_sched[qa + r] = rs;
      }
      h_tiles.push_back(H);
    };
    for (int64_t q = 1; q <= nrows; q++) {
      if (q == nrows || h_span[2 * q + 1] - h_span[2 * q0] > ACCW || q - q0 >= TILE_MAXROWS ||
          h_pairptr[q + 1] - h_pairptr[q0] > TILE_MAXPAIRS) {
        close_tile(q0, q);
        q0 = q;
      }
    }
  }
  if (too_many) PLAN_FAIL(NGSX_ERR_UNSUPPORTED, "rowplan_create: a dof with more than 255 adjacent list entries; use the coloured scatter");
  if (dcount != npairs) PLAN_FAIL(NGSX_ERR_CUDA, "rowplan_create: internal error (schedule does not cover the pairs)");
  P->ntiles = (int64_t)h_tiles.size();
  P->max_steps = max_steps;
  std::vector<int64_t> h_tile_d0(P->ntiles);
  for (int64_t c = 0; c < P->ntiles; c++) h_tile_d0[c] = h_tiles[c].d0;

  PLAN_CUDA(cudaMalloc(&d_tiles_nat, P->ntiles * sizeof(TileHdr)));
  PLAN_CUDA(cudaMalloc(&P->tiles, P->ntiles * sizeof(TileHdr)));
  PLAN_CUDA(cudaMalloc(&d_tile_d0, P->ntiles * sizeof(int64_t)));
  PLAN_CUDA(cudaMalloc(&d_tile_so, P->ntiles * sizeof(int64_t)));
  PLAN_CUDA(cudaMalloc(&d_stepoffs, std::max<size_t>(1, h_stepoffs.size()) * sizeof(int32_t)));
  PLAN_CUDA(cudaMalloc(&d_sched, nrows * sizeof(RowSched)));
  PLAN_CUDA(cudaMalloc(&P->desc, (size_t)npairs * P->ds));
  PLAN_CUDA(cudaMemcpyAsync(d_tiles_nat, h_tiles.data(), P->ntiles * sizeof(TileHdr), cudaMemcpyHostToDevice, st));
  PLAN_CUDA(cudaMemcpyAsync(d_tile_d0, h_tile_d0.data(), P->ntiles * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  PLAN_CUDA(cudaMemcpyAsync(d_tile_so, h_tile_so.data(), P->ntiles * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  PLAN_CUDA(cudaMemcpyAsync(d_stepoffs, h_stepoffs.data(), h_stepoffs.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  PLAN_CUDA(cudaMemcpyAsync(d_sched, h_sched.data(), nrows * sizeof(RowSched), cudaMemcpyHostToDevice, st));
  PLAN_CUDA(cudaMemsetAsync(P->desc, 0, (size_t)npairs * P->ds, st));
  k_plan_desc<<<(unsigned)((npairs + 127) / 128), 128, 0, st>>>(npairs, nrows, nd, nlist, P->ds, sorted, d_pairptr, d_sched, d_tile_d0, d_tile_so,
                                                               d_stepoffs, table, rowptr, colidx, P->desc, d_bad, P->lower);
  // tile processing order
  PLAN_CUDA(cudaMalloc(&d_ck, P->ntiles * sizeof(uint32_t)));
  PLAN_CUDA(cudaMalloc(&d_ck2, P->ntiles * sizeof(uint32_t)));
  PLAN_CUDA(cudaMalloc(&d_ids, P->ntiles * sizeof(int32_t)));
  PLAN_CUDA(cudaMalloc(&d_order, P->ntiles * sizeof(int32_t)));
  const int by_cost = nlist * 128 <= NGSX_ROWS_COSTORDER_BYTES;
  k_tile_keys<<<(unsigned)((P->ntiles + 255) / 256), 256, 0, st>>>(P->ntiles, d_tiles_nat, P->desc, P->ds, by_cost, d_ck, d_ids);
  cudaFree(tmp); tmp = nullptr; tmp_bytes = 0;
  PLAN_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_ck, d_ck2, d_ids, d_order, (int)P->ntiles, 0, 32, st));
  PLAN_CUDA(cudaMalloc(&tmp, tmp_bytes));
  PLAN_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, d_ck, d_ck2, d_ids, d_order, (int)P->ntiles, 0, 32, st));
  k_tile_permute<<<(unsigned)((P->ntiles + 255) / 256), 256, 0, st>>>(P->ntiles, d_order, d_tiles_nat, P->tiles);
  g_launches.fetch_add(5);
  int bad = 0;
  PLAN_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof bad, cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaStreamSynchronize(st));          // the host vectors go out of use only after the copies
  PLAN_CUDA(cudaGetLastError());
  if (bad & 1)
    PLAN_FAIL(NGSX_ERR_ARG, "rowplan_create: an element coupling is missing from the sparsity pattern (element_restriction too small?)");
  if (bad & 2)
    PLAN_FAIL(NGSX_ERR_UNSUPPORTED, "rowplan_create: a CSR row is longer than 256 entries; use the coloured scatter (ngsx_assemble_csr_h1)");
#undef PLAN_CUDA
#undef PLAN_FAIL
  cleanup();
  return NGSX_OK;
}

}  // extern "C"
