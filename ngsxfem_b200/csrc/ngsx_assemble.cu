// ngsx_assemble.cu -- (c) cut element matrices / vectors and (d) the fused colour-ordered scatter into CSR.
//
// Execution model (B200, 148 SMs): persistent CTAs of 8 warps; ONE WARP PER GROUP of consecutive list entries.
// A warp walks its group element by element; for each element all 32 lanes cooperate:
//   * every lane redundantly loads the 4 vertices / level-set values (one broadcast transaction) and derives the
//     affine geometry, the sign pattern and -- for cut elements -- the sub-simplex decomposition (table driven);
//   * uncut elements of the requested domain take the constant-rule path: the reference rule is the same on every
//     such element, so  sum_q w_q dphi_i/dxi_a dphi_j/dxi_b  is a per-CTA shared-memory table contracted with
//     J^-1 J^-T (same quadrature as the reference, re-associated);
//   * cut elements stage shape values/gradients of 16 quadrature points at a time in shared memory (lane q computes
//     point q) and every lane accumulates its 1..4 matrix entries over the staged points;
//   * sink: either a coalesced store of the dense nd x nd block, or "+=" into the CSR value array through the
//     precomputed slot map (coalesced int32 reads).  Groups of one colour share no dof and a warp processes its own
//     group sequentially, so there are no floating-point atomics and the result is bitwise reproducible.
// HBM-bound: 184 B in + 1600 B CSR read-modify-write per P2 element.
#include <algorithm>
#include <cstring>
#include <vector>
#include "ngsx_host.cuh"
#include "ngsx_rows.cuh"

namespace ngsx {

constexpr int WARPS = 8;          // warps per CTA
constexpr int QC = 16;            // quadrature points staged per round

struct FormDev { double cm, cs, cn, cd; int pm, ps, pn, pd; };

struct ElmatArgs {
  ngsx_mesh M;
  const double *lset;
  const CutTable *tab;
  int dt;
  const double4 *rule_vol; int nr_vol;
  const double4 *rule_if;  int nr_if;
  FormDev form;
  const int32_t *els;          // list of element numbers (NULL: identity)
  const int32_t *sublist;      // optional indirection: work item k -> list index sublist[k] (output stays at k)
  const int32_t *d_count;      // optional device-resident number of work items (overrides `count`)
  int group;                   // list entries per warp-sequential group
  int use_fast;                // constant-rule path for uncut elements
  const double *cf_vec; const int32_t *cf_el2dof; int cf_order;   // optional GridFunction factor of the whole form (H1 order 1-2): quadrature on every element
  int32_t *work; int32_t *work_count;   // order 3, dense sink: pass 1 records the quadrature-path entries here, pass 2 walks them (output at the entry)
  // sinks
  double *elmats; int32_t *npts;          // dense
  const int32_t *slots; double *vals;     // CSR
};

template <int DIM, int ORDER> struct Smem {
  static constexpr int ND = H1<DIM, ORDER>::ND;
  static constexpr int NE = ND * ND;
  static constexpr int NPAIR = DIM * (DIM + 1) / 2;
  // offsets in bytes
  static __host__ __device__ size_t rule_vol() { return 0; }
  static __host__ __device__ size_t rule_if(int nr_vol) { return (size_t)nr_vol * 32; }
  static __host__ __device__ size_t tab(int nr_vol, int nr_if) { return (size_t)(nr_vol + nr_if) * 32; }
  static __host__ __device__ size_t lut(int nr_vol, int nr_if) { return tab(nr_vol, nr_if) + (size_t)(NPAIR + 1) * NE * 8; }
  static __host__ __device__ size_t stage(int nr_vol, int nr_if) { return lut(nr_vol, nr_if) + 256; }
  static __host__ __device__ size_t stage_per_warp() { return (size_t)QC * ND * 32 + QC * 8; }
  static __host__ __device__ size_t total(int nr_vol, int nr_if) { return stage(nr_vol, nr_if) + WARPS * stage_per_warp(); }
};

template <int DIM>
__device__ __forceinline__ void load_el(const ngsx_mesh &M, const double *__restrict__ lset, int64_t e, double *ls, double (*p)[DIM], int *v_out = nullptr) {
  int v[4];
  if constexpr (DIM == 3) {
    int4 q = __ldg(reinterpret_cast<const int4 *>(M.el2v) + e);
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  } else {
    const int32_t *q = M.el2v + 3 * e;
    v[0] = __ldg(q); v[1] = __ldg(q + 1); v[2] = __ldg(q + 2);
  }
#pragma unroll
  for (int k = 0; k <= DIM; k++) {
    ls[k] = __ldg(lset + v[k]);
#pragma unroll
    for (int d = 0; d < DIM; d++) p[k][d] = __ldg(M.vert + (int64_t)v[k] * DIM + d);
    if (v_out) v_out[k] = v[k];
  }
}

template <int DIM>
__device__ __forceinline__ void bary_gradients(const ElGeom<DIM> &G, double (*gl)[DIM]) {
#pragma unroll
  for (int d = 0; d < DIM; d++) {
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; k++) { gl[k][d] = G.Jinv[k][d]; s += G.Jinv[k][d]; }
    gl[DIM][d] = -s;
  }
}

// SINK: 0 dense element matrices, 1 CSR read-modify-write
template <int DIM, int ORDER, int SINK>
__global__ void __launch_bounds__(WARPS * 32) k_elmat(ElmatArgs A, int64_t first, int64_t count) {
  using S = Smem<DIM, ORDER>;
  // order 3: the block is symmetric and never takes the constant-rule path -> lanes hold the lower triangle only (210 instead of 400
  // entries of a P3 tet: 7 instead of 13 accumulators per lane), mirrored at the sink
  constexpr bool SYM = ORDER == 3;
  constexpr int ND = S::ND, NE = S::NE, NPAIR = S::NPAIR, NENT = SYM ? ND * (ND + 1) / 2 : NE, T = (NENT + 31) / 32;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double4 *srule_vol = reinterpret_cast<double4 *>(smem_raw + S::rule_vol());
  double4 *srule_if = reinterpret_cast<double4 *>(smem_raw + S::rule_if(A.nr_vol));
  double *tab = reinterpret_cast<double *>(smem_raw + S::tab(A.nr_vol, A.nr_if));
  CutEntry *lut = reinterpret_cast<CutEntry *>(smem_raw + S::lut(A.nr_vol, A.nr_if));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char *wbase = smem_raw + S::stage(A.nr_vol, A.nr_if) + warp * S::stage_per_warp();
  double4 *stage = reinterpret_cast<double4 *>(wbase);
  double *facq = reinterpret_cast<double *>(wbase + (size_t)QC * ND * 32);

  // ---- CTA prologue: rules, decomposition table, constant-rule tables --------------------------------------
  for (int i = threadIdx.x; i < A.nr_vol; i += blockDim.x) srule_vol[i] = A.rule_vol[i];
  for (int i = threadIdx.x; i < A.nr_if; i += blockDim.x) srule_if[i] = A.rule_if[i];
  if (threadIdx.x < 16) lut[threadIdx.x] = A.tab->e[DIM - 2][A.dt][threadIdx.x];
  __syncthreads();
  const bool fast = A.use_fast && A.dt != IF && ORDER < 3 && !A.cf_vec;      // order 3: the edge functions depend on the element's vertex numbers
  if (fast) {
    double glr[DIM + 1][DIM];                      // reference gradients of the barycentrics
#pragma unroll
    for (int k = 0; k <= DIM; k++)
#pragma unroll
      for (int d = 0; d < DIM; d++) glr[k][d] = (k == DIM) ? -1.0 : (k == d ? 1.0 : 0.0);
    for (int idx = threadIdx.x; idx < NE; idx += blockDim.x) {
      int i = idx / ND, j = idx % ND;
      double acc[NPAIR + 1];
#pragma unroll
      for (int k = 0; k <= NPAIR; k++) acc[k] = 0.0;
      for (int q = 0; q < A.nr_vol; q++) {
        double4 r = srule_vol[q];
        double lam[4] = {r.x, r.y, r.z, 0.0};
        double last = 1.0;
#pragma unroll
        for (int k = 0; k < DIM; k++) last -= lam[k];
        lam[DIM] = last;
        double si, sj, gi[DIM], gj[DIM];
        h1_shape_grad<DIM, ORDER>(i, lam, glr, si, gi);
        h1_shape_grad<DIM, ORDER>(j, lam, glr, sj, gj);
        int pr = 0;
#pragma unroll
        for (int a = 0; a < DIM; a++)
#pragma unroll
          for (int b = a; b < DIM; b++) {
            acc[pr] += r.w * (a == b ? gi[a] * gj[a] : (gi[a] * gj[b] + gi[b] * gj[a]));
            pr++;
          }
        acc[NPAIR] += r.w * si * sj;
      }
#pragma unroll
      for (int k = 0; k <= NPAIR; k++) tab[k * NE + idx] = acc[k];
    }
    __syncthreads();
  }

  // per-lane matrix entries
  int ei[T], ej[T];
#pragma unroll
  for (int t = 0; t < T; t++) {
    int idx = lane + 32 * t;
    if constexpr (SYM) {
      int i = 0;
      while ((i + 1) * (i + 2) / 2 <= idx) i++;
      ei[t] = idx < NENT ? i : 0;
      ej[t] = idx < NENT ? idx - i * (i + 1) / 2 : 0;
    } else {
      ei[t] = idx < NE ? idx / ND : 0;
      ej[t] = idx < NE ? idx % ND : 0;
    }
  }

  if (A.d_count) count = min(count, (int64_t)*A.d_count);      // work list on the device, clamped to the caller's capacity
  const int64_t ngroups = (count + A.group - 1) / A.group;
  const int64_t nwarps = (int64_t)gridDim.x * WARPS;
  for (int64_t g = (int64_t)blockIdx.x * WARPS + warp; g < ngroups; g += nwarps) {
    const int64_t kb = first + g * A.group;
    const int64_t ke = min(kb + (int64_t)A.group, first + count);
    for (int64_t k = kb; k < ke; k++) {
      const int64_t kk = A.sublist ? A.sublist[k] : k;
      const int64_t e = A.els ? A.els[kk] : kk;
      double ls[4], p[DIM + 1][DIM];
      int vn[4];
      load_el<DIM>(A.M, A.lset, e, ls, p, vn);
      const unsigned orient = ORDER == 3 ? edge_orientation<DIM>(vn) : 0u;
      int mask;
      const int dom = sign_test<DIM>(ls, mask);
      const bool cut = dom == IF;
      double acc[T];
#pragma unroll
      for (int t = 0; t < T; t++) acc[t] = 0.0;
      int np_used = 0;
      const bool active = cut || dom == A.dt;
      if (active) {
        ElGeom<DIM> G;
        element_geometry<DIM>(p, G);
        const double meas = fabs(G.det);
        double cm = A.form.cm, cs = A.form.cs, cn = A.form.cn, cd = A.form.cd;
        if (A.form.pm | A.form.ps | A.form.pn | A.form.pd) {
          double h = DIM == 2 ? sqrt(meas) : cbrt(meas);           // specialcf.mesh_size = |det J|^(1/dim)
          cm *= ipow(h, A.form.pm); cs *= ipow(h, A.form.ps); cn *= ipow(h, A.form.pn); cd *= ipow(h, A.form.pd);
        }
        if (!cut && fast) {
          // ---- constant-rule path -------------------------------------------------------------------------
          double gs[NPAIR];
          int pr = 0;
#pragma unroll
          for (int a = 0; a < DIM; a++)
#pragma unroll
            for (int b = a; b < DIM; b++) {
              double s = 0.0;
#pragma unroll
              for (int d = 0; d < DIM; d++) s += G.Jinv[a][d] * G.Jinv[b][d];
              gs[pr++] = s;
            }
#pragma unroll
          for (int t = 0; t < T; t++) {
            int idx = lane + 32 * t;
            if (!SYM && idx < NE) {
              double s = 0.0;
#pragma unroll
              for (int q = 0; q < NPAIR; q++) s += gs[q] * tab[q * NE + idx];
              acc[t] = meas * (cs * s + cm * tab[NPAIR * NE + idx]);
            }
          }
          np_used = A.nr_vol;
        } else {
          // ---- quadrature path ----------------------------------------------------------------------------
          double gl[DIM + 1][DIM];
          bary_gradients<DIM>(G, gl);
          double nrm[DIM];
          double scale = 1.0;
          {
            double g2 = 0.0, r2 = 0.0;
#pragma unroll
            for (int d = 0; d < DIM; d++) {
              double v = 0.0;
#pragma unroll
              for (int kk = 0; kk <= DIM; kk++) v += ls[kk] * gl[kk][d];
              nrm[d] = v; g2 += v * v;
            }
#pragma unroll
            for (int kk = 0; kk < DIM; kk++) { double v = ls[kk] - ls[DIM]; r2 += v * v; }
            double inv = 1.0 / sqrt(g2);
#pragma unroll
            for (int d = 0; d < DIM; d++) nrm[d] *= inv;
            if (cut && A.dt == IF) scale = sqrt(g2) / sqrt(r2);       // |F^-T n_ref|
          }
          CutEntry E;
          if (cut) E = lut[mask]; else { E.ns = 1; E.npt = 0; }
          const double4 *rule = (cut && A.dt == IF) ? srule_if : srule_vol;
          const int nr = (cut && A.dt == IF) ? A.nr_if : A.nr_vol;
          np_used = E.ns * nr;
          for (int s = 0; s < E.ns; s++) {
            double P[4][3];
            double sfac = 1.0;
            if (cut) {
              for (int v = 0; v < E.npt; v++) coded_point<DIM>(E.v[s][v], ls, P[v]);
              sfac = simplex_factor(E.npt, P);
            }
            for (int q0 = 0; q0 < nr; q0 += QC) {
              const int nq = min(QC, nr - q0);
              if (lane < nq) {
                double4 r = rule[q0 + lane];
                double xi[3] = {r.x, r.y, r.z};
                if (cut) simplex_point(E.npt, P, r, xi);
                double lam[4] = {xi[0], xi[1], xi[2], 0.0};
                double last = 1.0;
#pragma unroll
                for (int kk = 0; kk < DIM; kk++) last -= lam[kk];
                lam[DIM] = last;
#pragma unroll
                for (int i = 0; i < ND; i++) {
                  double sv, gv[DIM];
                  h1_shape_grad<DIM, ORDER>(i, lam, gl, sv, gv, orient);
                  stage[lane * ND + i] = make_double4(sv, gv[0], gv[1], DIM == 3 ? gv[DIM - 1] : 0.0);
                }
                double gfv = 1.0;
                if (A.cf_vec) {                                  // GridFunction factor at the point: sum_i c_i phi_i (orientation-free orders 1, 2)
                  gfv = 0.0;
                  const int32_t *cd = A.cf_el2dof + e * (A.cf_order == 1 ? DIM + 1 : H1<DIM, 2>::ND);
                  if (A.cf_order == 1) {
#pragma unroll
                    for (int i = 0; i <= DIM; i++) gfv += __ldg(A.cf_vec + __ldg(cd + i)) * h1_shape<DIM, 1>(i, lam);
                  } else {
#pragma unroll
                    for (int i = 0; i < H1<DIM, 2>::ND; i++) gfv += __ldg(A.cf_vec + __ldg(cd + i)) * h1_shape<DIM, 2>(i, lam);
                  }
                }
                facq[lane] = meas * (r.w * sfac * scale) * gfv;
              }
              __syncwarp();
              for (int q = 0; q < nq; q++) {
                const double f = facq[q];
#pragma unroll
                for (int t = 0; t < T; t++) {
                  const double4 bi = stage[q * ND + ei[t]], bj = stage[q * ND + ej[t]];
                  double gg = bi.y * bj.y + bi.z * bj.z + (DIM == 3 ? bi.w * bj.w : 0.0);
                  double v = cs * gg + cm * bi.x * bj.x;
                  if (cn != 0.0 || cd != 0.0) {
                    double gin = bi.y * nrm[0] + bi.z * nrm[1] + (DIM == 3 ? bi.w * nrm[DIM - 1] : 0.0);
                    double gjn = bj.y * nrm[0] + bj.z * nrm[1] + (DIM == 3 ? bj.w * nrm[DIM - 1] : 0.0);
                    v += cn * (gjn * bi.x + gin * bj.x) + cd * gin * gjn;
                  }
                  acc[t] += f * v;
                }
              }
              __syncwarp();
            }
          }
        }
      }
      // ---- sink ---------------------------------------------------------------------------------------------
      if constexpr (SINK == 0) {
#pragma unroll
        for (int t = 0; t < T; t++) {
          int idx = lane + 32 * t;
          if constexpr (SYM) {
            if (idx < NENT) {
              A.elmats[k * NE + ei[t] * ND + ej[t]] = acc[t];
              if (ei[t] != ej[t]) A.elmats[k * NE + ej[t] * ND + ei[t]] = acc[t];
            }
          } else {
            if (idx < NE) A.elmats[k * NE + idx] = acc[t];
          }
        }
        if (A.npts && lane == 0) A.npts[k] = np_used;
      } else {
        if (active) {
#pragma unroll
          for (int t = 0; t < T; t++) {
            int idx = lane + 32 * t;
            if constexpr (SYM) {
              if (idx < NENT) {                               // (i, j) and (j, i) are distinct CSR entries: no two lanes share a slot
                A.vals[__ldg(A.slots + k * NE + ei[t] * ND + ej[t])] += acc[t];
                if (ei[t] != ej[t]) A.vals[__ldg(A.slots + k * NE + ej[t] * ND + ei[t])] += acc[t];
              }
            } else if (idx < NE) {
              int32_t sl = __ldg(A.slots + k * NE + idx);
              A.vals[sl] += acc[t];
            }
          }
        }
        __syncwarp();          // next element of the group may touch the same CSR entries from another lane
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// H1 order 3 element matrices (row g-1): warp per element, W3 warps per CTA.
//  * uncut elements of the requested domain: constant-rule tables after all.  The only orientation dependence of the order-3
//    basis is the SIGN of the second function of each edge (x = +-(lam_b - lam_a), NGS/fem/h1hofe_impl.hpp:78-108), so the block
//    is  S T(J) S  with S = diag(+-1) and T the table contraction of the orientation-free element: 7 FMAs per entry instead of
//    24 points x 4, no staging, one coalesced store of the block.
//  * cut elements (and IF / Nitsche forms): the quadrature contraction  sum_q f_q b_i(q)^T Q b_j(q),  b = (phi, grad phi) in R^4,
//    Q the 4x4 symmetric matrix of the form ([[c_m, c_n n^T],[c_n n, c_s I + c_d n n^T]]), runs on the FP64 tensor cores:
//    per staged point ONE k-step of mma.sync.m8n8k4.f64 (DMMA.884) for each lower 8x8 tile of the block padded to 24 x 24
//    (6 tiles; 3 for a P3 triangle).  A fragment = component lane%4 of b_{8r + lane/4}, B fragment = f_q (Q b_{8c + lane/4})[lane%4],
//    both from one 32-byte row read (4 lanes share a row: broadcast, no bank conflict) -- 3 LDS.256 + 6 DMMA per point where
//    the FMA version (k_elmat<DIM,3>) issued 14 LDS.256 with 5-way conflicts for its 7 entries per lane.
//    The finished fragments go through a per-warp block in shared memory, mirrored, and leave as one coalesced store.
// ---------------------------------------------------------------------------------------------------------------
constexpr int W3 = 4;             // warps per CTA of the order-3 kernel
constexpr int QC3 = 32;           // points staged per round (one per lane)

template <int DIM> struct Smem3 {
  static constexpr int ND = H1<DIM, 3>::ND, NE = ND * ND, NTRI = ND * (ND + 1) / 2, NPAIR = DIM * (DIM + 1) / 2;
  static constexpr int NT = (ND + 7) / 8, NDP = NT * 8;
  // double4 per staged point: ND rows + 1 (odd: the 32 writers of a row spread over all bank groups).  The last tile reads rows
  // ND..NDP-1, i.e. the next point's first rows or, behind the last point, the factor array: finite or not, those values reach
  // padded entries of the tiles only.
  static constexpr int PSTR4 = (ND | 1) == ND ? ND + 2 : ND + 1;
  static __host__ __device__ size_t tab(int nr_vol, int nr_if) { return (size_t)(nr_vol + nr_if) * 32; }
  static __host__ __device__ size_t lut(int nr_vol, int nr_if) { return tab(nr_vol, nr_if) + (size_t)(NPAIR + 1) * NTRI * 8; }
  static __host__ __device__ size_t stage(int nr_vol, int nr_if) { return (lut(nr_vol, nr_if) + 256 + 31) / 32 * 32; }
  static __host__ __device__ size_t stage_per_warp() {
    size_t a = (size_t)QC3 * PSTR4 * 32 + QC3 * 8 + 256, b = (size_t)NE * 8;       // the output block aliases the stage
    return ((a > b ? a : b) + 31) / 32 * 32;
  }
  static __host__ __device__ size_t total(int nr_vol, int nr_if) { return stage(nr_vol, nr_if) + W3 * stage_per_warp(); }
};

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// dofs whose sign flips with the edge orientation: the second function of edge e (local dof DIM + 1 + 2 e + 1)
template <int DIM> __device__ __forceinline__ unsigned flipped_dofs(unsigned orient) {
  unsigned m = 0;
#pragma unroll
  for (int e = 0; e < H1<DIM, 3>::NEDGE; e++) m |= ((orient >> e) & 1u) << (DIM + 2 + 2 * e);
  return m;
}

// tables of the orientation-free order-3 element, lower triangle: tab[q][tri(i, j)] = sum_p w_p d_a phi_i d_b phi_j (pair q = (a, b),
// reference gradients), tab[NPAIR][.] = sum_p w_p phi_i phi_j.  All threads of the CTA.
template <int DIM>
__device__ __forceinline__ void order3_tables(double *tab, const double4 *srule_vol, int nr_vol) {
  constexpr int ND = H1<DIM, 3>::ND, NTRI = ND * (ND + 1) / 2, NPAIR = DIM * (DIM + 1) / 2;
  double glr[DIM + 1][DIM];
#pragma unroll
  for (int k = 0; k <= DIM; k++)
#pragma unroll
    for (int d = 0; d < DIM; d++) glr[k][d] = (k == DIM) ? -1.0 : (k == d ? 1.0 : 0.0);
  for (int idx = threadIdx.x; idx < NTRI; idx += blockDim.x) {
    int i = 0;
    while ((i + 1) * (i + 2) / 2 <= idx) i++;
    const int j = idx - i * (i + 1) / 2;
    double acc[NPAIR + 1];
#pragma unroll
    for (int k = 0; k <= NPAIR; k++) acc[k] = 0.0;
    for (int q = 0; q < nr_vol; q++) {
      double4 r = srule_vol[q];
      double lam[4] = {r.x, r.y, r.z, 0.0};
      double last = 1.0;
#pragma unroll
      for (int k = 0; k < DIM; k++) last -= lam[k];
      lam[DIM] = last;
      double si, sj, gi[DIM], gj[DIM];
      h1_shape_grad<DIM, 3>(i, lam, glr, si, gi, 0u);
      h1_shape_grad<DIM, 3>(j, lam, glr, sj, gj, 0u);
      int pr = 0;
#pragma unroll
      for (int a = 0; a < DIM; a++)
#pragma unroll
        for (int b = a; b < DIM; b++) {
          acc[pr] += r.w * (a == b ? gi[a] * gj[a] : (gi[a] * gj[b] + gi[b] * gj[a]));
          pr++;
        }
      acc[NPAIR] += r.w * si * sj;
    }
#pragma unroll
    for (int k = 0; k <= NPAIR; k++) tab[k * NTRI + idx] = acc[k];
  }
}

// pass 1 of the dense sink: warp per PAIR of consecutive list entries (every table value read from shared memory serves two
// blocks).  Uncut entries of the requested domain: block from the tables (see above), one coalesced store.  Entries outside the
// domain: zeros (same code, zero factors).  Cut entries: recorded in the work list for k_elmat3 (integer slot from an atomic counter:
// the order of the list never reaches a value, every block is written at its own entry).  The tables are kept as full nd x nd
// arrays here: lane l reads entry l + 32 t, conflict free.
constexpr int W3F = 8;
template <int DIM>
__global__ void __launch_bounds__(W3F * 32, 3) k_elmat3_fast(ElmatArgs A, int64_t count) {
  constexpr int ND = H1<DIM, 3>::ND, NE = ND * ND, NTRI = ND * (ND + 1) / 2, NPAIR = DIM * (DIM + 1) / 2, T = (NE + 31) / 32;
  extern __shared__ __align__(32) unsigned char smem_raw[];
  double4 *srule_vol = reinterpret_cast<double4 *>(smem_raw);
  double *tab = reinterpret_cast<double *>(smem_raw + (size_t)A.nr_vol * 32);          // (NPAIR + 1) x NE
  double *tri = tab + (NPAIR + 1) * NE;                                                // (NPAIR + 1) x NTRI, set-up only
  for (int i = threadIdx.x; i < A.nr_vol; i += blockDim.x) srule_vol[i] = A.rule_vol[i];
  __syncthreads();
  order3_tables<DIM>(tri, srule_vol, A.nr_vol);
  __syncthreads();
  for (int idx = threadIdx.x; idx < NE; idx += blockDim.x) {
    const int i = idx / ND, j = idx % ND, a = max(i, j), b = min(i, j);
#pragma unroll
    for (int q = 0; q <= NPAIR; q++) tab[q * NE + idx] = tri[q * NTRI + a * (a + 1) / 2 + b];
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned short eij[T];
#pragma unroll
  for (int t = 0; t < T; t++) {
    const int idx = min(lane + 32 * t, NE - 1);
    eij[t] = (unsigned short)((idx / ND) | ((idx % ND) << 8));
  }
  const int64_t npair = (count + 1) / 2;
  const int64_t nwarps = (int64_t)gridDim.x * W3F;
  for (int64_t kp = (int64_t)blockIdx.x * W3F + warp; kp < npair; kp += nwarps) {
    double gs[2][NPAIR + 1];
    unsigned flip[2];
    bool store[2];
    int np[2];
#pragma unroll
    for (int u = 0; u < 2; u++) {
      const int64_t k = 2 * kp + u;
      store[u] = false; flip[u] = 0u; np[u] = 0;
#pragma unroll
      for (int q = 0; q <= NPAIR; q++) gs[u][q] = 0.0;
      if (k >= count) continue;
      const int64_t e = A.els ? A.els[k] : k;
      double ls[4], p[DIM + 1][DIM];
      int vn[4];
      load_el<DIM>(A.M, A.lset, e, ls, p, vn);
      int mask;
      const int dom = sign_test<DIM>(ls, mask);
      if (dom == IF) {
        if (lane == 0) A.work[atomicAdd(A.work_count, 1)] = (int32_t)k;
        continue;
      }
      store[u] = true;
      if (dom != A.dt) continue;
      np[u] = A.nr_vol;
      ElGeom<DIM> G;
      element_geometry<DIM>(p, G);
      const double meas = fabs(G.det);
      double cm = A.form.cm, cs = A.form.cs;
      if (A.form.pm | A.form.ps) {
        double h = DIM == 2 ? sqrt(meas) : cbrt(meas);
        cm *= ipow(h, A.form.pm); cs *= ipow(h, A.form.ps);
      }
      int pr = 0;
#pragma unroll
      for (int a = 0; a < DIM; a++)
#pragma unroll
        for (int b = a; b < DIM; b++) {
          double s = 0.0;
#pragma unroll
          for (int d = 0; d < DIM; d++) s += G.Jinv[a][d] * G.Jinv[b][d];
          gs[u][pr++] = meas * cs * s;
        }
      gs[u][NPAIR] = meas * cm;
      flip[u] = flipped_dofs<DIM>(edge_orientation<DIM>(vn));
    }
    double *o0 = A.elmats + 2 * kp * NE, *o1 = o0 + NE;
#pragma unroll
    for (int t = 0; t < T; t++) {
      const int idx = lane + 32 * t;
      if (idx < NE) {
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int q = 0; q <= NPAIR; q++) {
          const double w = tab[q * NE + idx];
          s0 += gs[0][q] * w; s1 += gs[1][q] * w;
        }
        const int i = eij[t] & 255, j = eij[t] >> 8;
        if (((flip[0] >> i) ^ (flip[0] >> j)) & 1u) s0 = -s0;
        if (((flip[1] >> i) ^ (flip[1] >> j)) & 1u) s1 = -s1;
        if (store[0]) o0[idx] = s0;
        if (store[1]) o1[idx] = s1;
      }
    }
    if (A.npts && lane < 2 && store[lane]) A.npts[2 * kp + lane] = np[lane];
  }
}

template <int DIM, int SINK>
__global__ void __launch_bounds__(W3 * 32) k_elmat3(ElmatArgs A, int64_t first, int64_t count) {
  using S = Smem3<DIM>;
  constexpr int ND = S::ND, NE = S::NE, NTRI = S::NTRI, NPAIR = S::NPAIR, NT = S::NT, PSTR4 = S::PSTR4, T = (NE + 31) / 32;
  constexpr int NTILE = NT * (NT + 1) / 2;
  extern __shared__ __align__(32) unsigned char smem_raw[];
  double4 *srule_vol = reinterpret_cast<double4 *>(smem_raw);
  double4 *srule_if = srule_vol + A.nr_vol;
  double *tab = reinterpret_cast<double *>(smem_raw + S::tab(A.nr_vol, A.nr_if));
  CutEntry *lut = reinterpret_cast<CutEntry *>(smem_raw + S::lut(A.nr_vol, A.nr_if));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char *wbase = smem_raw + S::stage(A.nr_vol, A.nr_if) + warp * S::stage_per_warp();
  double4 *stage = reinterpret_cast<double4 *>(wbase);
  double *facq = reinterpret_cast<double *>(wbase + (size_t)QC3 * PSTR4 * 32);
  double *block = reinterpret_cast<double *>(wbase);               // aliases the stage (used after the contraction)

  for (int i = threadIdx.x; i < A.nr_vol; i += blockDim.x) srule_vol[i] = A.rule_vol[i];
  for (int i = threadIdx.x; i < A.nr_if; i += blockDim.x) srule_if[i] = A.rule_if[i];
  if (threadIdx.x < 16) lut[threadIdx.x] = A.tab->e[DIM - 2][A.dt][threadIdx.x];
  __syncthreads();
  const bool fast = A.use_fast && A.dt != IF;
  if (fast) {
    order3_tables<DIM>(tab, srule_vol, A.nr_vol);
    __syncthreads();
  }

  // full-block entries of this lane (output order) with their place in the lower-triangle tables
  int etri[T];
  unsigned short eij[T];
#pragma unroll
  for (int t = 0; t < T; t++) {
    const int idx = min(lane + 32 * t, NE - 1);
    const int i = idx / ND, j = idx % ND, a = max(i, j), b = min(i, j);
    etri[t] = a * (a + 1) / 2 + b;
    eij[t] = (unsigned short)(i | (j << 8));
  }
  const int kq = lane & 3, rq = lane >> 2;

  if (A.d_count) count = min(count, (int64_t)*A.d_count);
  if (A.work) count = *A.work_count;                             // pass 2 of the dense sink: the recorded entries (group = 1, first = 0)
  const int64_t ngroups = (count + A.group - 1) / A.group;
  const int64_t nwarps = (int64_t)gridDim.x * W3;
  for (int64_t g = (int64_t)blockIdx.x * W3 + warp; g < ngroups; g += nwarps) {
    const int64_t kb = first + g * A.group;
    const int64_t ke = min(kb + (int64_t)A.group, first + count);
    for (int64_t k0 = kb; k0 < ke; k0++) {
      const int64_t k = A.work ? A.work[k0] : k0;
      const int64_t kk = A.sublist ? A.sublist[k] : k;
      const int64_t e = A.els ? A.els[kk] : kk;
      double ls[4], p[DIM + 1][DIM];
      int vn[4];
      load_el<DIM>(A.M, A.lset, e, ls, p, vn);
      const unsigned orient = edge_orientation<DIM>(vn);
      int mask;
      const int dom = sign_test<DIM>(ls, mask);
      const bool cut = dom == IF;
      const bool active = cut || dom == A.dt;
      int np_used = 0;
      bool in_block = false;                 // result sits in the shared-memory block (quadrature path) or in out[] (table path / zero)
      double out[T];
#pragma unroll
      for (int t = 0; t < T; t++) out[t] = 0.0;
      if (active) {
        ElGeom<DIM> G;
        element_geometry<DIM>(p, G);
        const double meas = fabs(G.det);
        double cm = A.form.cm, cs = A.form.cs, cn = A.form.cn, cd = A.form.cd;
        if (A.form.pm | A.form.ps | A.form.pn | A.form.pd) {
          double h = DIM == 2 ? sqrt(meas) : cbrt(meas);
          cm *= ipow(h, A.form.pm); cs *= ipow(h, A.form.ps); cn *= ipow(h, A.form.pn); cd *= ipow(h, A.form.pd);
        }
        if (!cut && fast) {
          double gs[NPAIR];
          int pr = 0;
#pragma unroll
          for (int a = 0; a < DIM; a++)
#pragma unroll
            for (int b = a; b < DIM; b++) {
              double s = 0.0;
#pragma unroll
              for (int d = 0; d < DIM; d++) s += G.Jinv[a][d] * G.Jinv[b][d];
              gs[pr++] = meas * cs * s;
            }
          const double gmass = meas * cm;
          const unsigned flip = flipped_dofs<DIM>(orient);
#pragma unroll
          for (int t = 0; t < T; t++) {
            const int tr = etri[t];
            double s = gmass * tab[NPAIR * NTRI + tr];
#pragma unroll
            for (int q = 0; q < NPAIR; q++) s += gs[q] * tab[q * NTRI + tr];
            const unsigned sg = ((flip >> (eij[t] & 255)) ^ (flip >> (eij[t] >> 8))) & 1u;
            out[t] = sg ? -s : s;
          }
          np_used = A.nr_vol;
        } else {
          in_block = true;
          double gl[DIM + 1][DIM];
          bary_gradients<DIM>(G, gl);
          double nrm[3] = {0.0, 0.0, 0.0};
          double scale = 1.0;
          {
            double g2 = 0.0, r2 = 0.0;
#pragma unroll
            for (int d = 0; d < DIM; d++) {
              double v = 0.0;
#pragma unroll
              for (int m = 0; m <= DIM; m++) v += ls[m] * gl[m][d];
              nrm[d] = v; g2 += v * v;
            }
#pragma unroll
            for (int m = 0; m < DIM; m++) { double v = ls[m] - ls[DIM]; r2 += v * v; }
            double inv = 1.0 / sqrt(g2);
#pragma unroll
            for (int d = 0; d < DIM; d++) nrm[d] *= inv;
            if (cut && A.dt == IF) scale = sqrt(g2) / sqrt(r2);
          }
          // row kq of Q (diagonal case: Q0 = the lane's diagonal entry)
          const bool diagq = cn == 0.0 && cd == 0.0;
          double Q0, Q1, Q2, Q3;
          if (diagq) { Q0 = kq == 0 ? cm : cs; Q1 = Q2 = Q3 = 0.0; }
          else if (kq == 0) { Q0 = cm; Q1 = cn * nrm[0]; Q2 = cn * nrm[1]; Q3 = cn * nrm[2]; }
          else {
            const double nk = kq == 1 ? nrm[0] : (kq == 2 ? nrm[1] : nrm[2]);
            Q0 = cn * nk;
            Q1 = cd * nk * nrm[0] + (kq == 1 ? cs : 0.0);
            Q2 = cd * nk * nrm[1] + (kq == 2 ? cs : 0.0);
            Q3 = cd * nk * nrm[2] + (kq == 3 ? cs : 0.0);
          }
          double acc[NTILE][2];
#pragma unroll
          for (int t = 0; t < NTILE; t++) acc[t][0] = acc[t][1] = 0.0;
          CutEntry E;
          if (cut) E = lut[mask]; else { E.ns = 1; E.npt = 0; }
          const double4 *rule = (cut && A.dt == IF) ? srule_if : srule_vol;
          const int nr = (cut && A.dt == IF) ? A.nr_if : A.nr_vol;
          np_used = E.ns * nr;
          for (int s = 0; s < E.ns; s++) {
            double P[4][3];
            double sfac = 1.0;
            if (cut) {
              for (int v = 0; v < E.npt; v++) coded_point<DIM>(E.v[s][v], ls, P[v]);
              sfac = simplex_factor(E.npt, P);
            }
            for (int q0 = 0; q0 < nr; q0 += QC3) {
              const int nq = min(QC3, nr - q0);
              if (lane < nq) {
                double4 r = rule[q0 + lane];
                double xi[3] = {r.x, r.y, r.z};
                if (cut) simplex_point(E.npt, P, r, xi);
                double lam[4] = {xi[0], xi[1], xi[2], 0.0};
                double last = 1.0;
#pragma unroll
                for (int m = 0; m < DIM; m++) last -= lam[m];
                lam[DIM] = last;
#pragma unroll
                for (int i = 0; i < ND; i++) {
                  double sv, gv[DIM];
                  h1_shape_grad<DIM, 3>(i, lam, gl, sv, gv, orient);
                  stage[lane * PSTR4 + i] = make_double4(sv, gv[0], gv[1], DIM == 3 ? gv[DIM - 1] : 0.0);
                }
                facq[lane] = meas * (r.w * sfac * scale);
              }
              __syncwarp();
              const double *stage1 = reinterpret_cast<const double *>(stage) + rq * 4 + kq;      // component kq of row rq
              if (diagq) {
                // Q = diag(c_m, c_s, c_s, c_s): B fragment = A fragment x f_q Q_kk, no second read
#pragma unroll 2
                for (int q = 0; q < nq; q++) {
                  const double fd = facq[q] * Q0;
                  double af[NT], bf[NT];
#pragma unroll
                  for (int c = 0; c < NT; c++) { af[c] = stage1[(q * PSTR4 + 8 * c) * 4]; bf[c] = af[c] * fd; }
                  int t = 0;
#pragma unroll
                  for (int r = 0; r < NT; r++)
#pragma unroll
                    for (int c = 0; c <= r; c++) { dmma884(acc[t][0], acc[t][1], af[r], bf[c]); t++; }
                }
              } else {
                for (int q = 0; q < nq; q++) {
                  const double f = facq[q];
                  double af[NT], bf[NT];
#pragma unroll
                  for (int c = 0; c < NT; c++) {
                    const double4 v = stage[q * PSTR4 + 8 * c + rq];
                    af[c] = stage1[(q * PSTR4 + 8 * c) * 4];
                    bf[c] = f * (Q0 * v.x + Q1 * v.y + Q2 * v.z + Q3 * v.w);
                  }
                  int t = 0;
#pragma unroll
                  for (int r = 0; r < NT; r++)
#pragma unroll
                    for (int c = 0; c <= r; c++) { dmma884(acc[t][0], acc[t][1], af[r], bf[c]); t++; }
                }
              }
              __syncwarp();
            }
          }
          // fragments -> block (mirrored).  lane holds C[8r + rq][8c + 2kq + {0,1}]
          {
            int t = 0;
#pragma unroll
            for (int r = 0; r < NT; r++)
#pragma unroll
              for (int c = 0; c <= r; c++) {
                const int i = 8 * r + rq;
#pragma unroll
                for (int h = 0; h < 2; h++) {
                  const int j = 8 * c + 2 * kq + h;
                  if (i < ND && j < ND) {
                    block[i * ND + j] = acc[t][h];
                    if (r != c) block[j * ND + i] = acc[t][h];
                  }
                }
                t++;
              }
          }
          __syncwarp();
        }
      }
      // ---- sink ---------------------------------------------------------------------------------------------
#pragma unroll
      for (int t = 0; t < T; t++) {
        const int idx = lane + 32 * t;
        if (idx < NE) {
          const double v = in_block ? block[idx] : out[t];
          if constexpr (SINK == 0) A.elmats[k * NE + idx] = v;
          else if (active) A.vals[__ldg(A.slots + k * NE + idx)] += v;
        }
      }
      if (SINK == 0 && A.npts && lane == 0) A.npts[k] = np_used;
      __syncwarp();          // the block aliases the next element's stage; coloured sink: the next element of the group may share entries
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// element vectors (SymbolicCutLFI): same staging, lanes < ND accumulate  sum_q fac_q f(x_q) phi_i(x_q)
// SINK 0: dense elvecs[k][ND]; 1: vec[el2dof_list[k][i]] += (coloured groups)
// ---------------------------------------------------------------------------------------------------------------
struct PolyDev2 { int n; double c[NGSX_POLY_MAXTERMS]; int ex[NGSX_POLY_MAXTERMS][3]; };

struct ElvecArgs {
  ngsx_mesh M;
  const double *lset;
  const CutTable *tab;
  int dt;
  const double4 *rule_vol; int nr_vol;
  const double4 *rule_if;  int nr_if;
  PolyDev2 f;
  const int32_t *els;
  int group;
  double *elvecs; int32_t *npts;
  const int32_t *el2dof_list; double *vec;
};

template <int DIM, int ORDER, int SINK>
__global__ void __launch_bounds__(WARPS * 32) k_elvec(ElvecArgs A, int64_t first, int64_t count) {
  constexpr int ND = H1<DIM, ORDER>::ND;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double4 *srule_vol = reinterpret_cast<double4 *>(smem_raw);
  double4 *srule_if = srule_vol + A.nr_vol;
  CutEntry *lut = reinterpret_cast<CutEntry *>(srule_if + A.nr_if);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *stage = reinterpret_cast<double *>(smem_raw + (size_t)(A.nr_vol + A.nr_if) * 32 + 256) + warp * (32 * ND + 32);
  double *facq = stage + 32 * ND;
  for (int i = threadIdx.x; i < A.nr_vol; i += blockDim.x) srule_vol[i] = A.rule_vol[i];
  for (int i = threadIdx.x; i < A.nr_if; i += blockDim.x) srule_if[i] = A.rule_if[i];
  if (threadIdx.x < 16) lut[threadIdx.x] = A.tab->e[DIM - 2][A.dt][threadIdx.x];
  __syncthreads();
  const int64_t ngroups = (count + A.group - 1) / A.group;
  const int64_t nwarps = (int64_t)gridDim.x * WARPS;
  for (int64_t g = (int64_t)blockIdx.x * WARPS + warp; g < ngroups; g += nwarps) {
    const int64_t kb = first + g * A.group;
    const int64_t ke = min(kb + (int64_t)A.group, first + count);
    for (int64_t k = kb; k < ke; k++) {
      const int64_t e = A.els ? A.els[k] : k;
      double ls[4], p[DIM + 1][DIM];
      int vn[4];
      load_el<DIM>(A.M, A.lset, e, ls, p, vn);
      const unsigned orient = ORDER == 3 ? edge_orientation<DIM>(vn) : 0u;
      int mask;
      const int dom = sign_test<DIM>(ls, mask);
      const bool cut = dom == IF;
      const bool active = cut || dom == A.dt;
      double acc = 0.0;
      int np_used = 0;
      if (active) {
        ElGeom<DIM> G;
        element_geometry<DIM>(p, G);
        const double meas = fabs(G.det);
        double scale = 1.0;
        if (cut && A.dt == IF) {
          double gl[DIM + 1][DIM];
          bary_gradients<DIM>(G, gl);
          double g2 = 0.0, r2 = 0.0;
#pragma unroll
          for (int d = 0; d < DIM; d++) {
            double v = 0.0;
#pragma unroll
            for (int kk = 0; kk <= DIM; kk++) v += ls[kk] * gl[kk][d];
            g2 += v * v;
          }
#pragma unroll
          for (int kk = 0; kk < DIM; kk++) { double v = ls[kk] - ls[DIM]; r2 += v * v; }
          scale = sqrt(g2) / sqrt(r2);
        }
        CutEntry E;
        if (cut) E = lut[mask]; else { E.ns = 1; E.npt = 0; }
        const double4 *rule = (cut && A.dt == IF) ? srule_if : srule_vol;
        const int nr = (cut && A.dt == IF) ? A.nr_if : A.nr_vol;
        np_used = E.ns * nr;
        for (int s = 0; s < E.ns; s++) {
          double P[4][3];
          double sfac = 1.0;
          if (cut) {
            for (int v = 0; v < E.npt; v++) coded_point<DIM>(E.v[s][v], ls, P[v]);
            sfac = simplex_factor(E.npt, P);
          }
          for (int q0 = 0; q0 < nr; q0 += 32) {
            const int nq = min(32, nr - q0);
            if (lane < nq) {
              double4 r = rule[q0 + lane];
              double xi[3] = {r.x, r.y, r.z};
              if (cut) simplex_point(E.npt, P, r, xi);
              double lam[4] = {xi[0], xi[1], xi[2], 0.0};
              double last = 1.0;
#pragma unroll
              for (int kk = 0; kk < DIM; kk++) last -= lam[kk];
              lam[DIM] = last;
              double x[3] = {0.0, 0.0, 0.0};
#pragma unroll
              for (int d = 0; d < DIM; d++) {
                double v = 0.0;
#pragma unroll
                for (int kk = 0; kk <= DIM; kk++) v += lam[kk] * p[kk][d];
                x[d] = v;
              }
              double fx = 0.0;
              for (int m = 0; m < A.f.n; m++)
                fx += A.f.c[m] * ipow(x[0], A.f.ex[m][0]) * ipow(x[1], A.f.ex[m][1]) * ipow(x[2], A.f.ex[m][2]);
#pragma unroll
              for (int i = 0; i < ND; i++) stage[lane * ND + i] = h1_shape<DIM, ORDER>(i, lam, orient);
              facq[lane] = meas * (r.w * sfac * scale) * fx;
            }
            __syncwarp();
            if (lane < ND)
              for (int q = 0; q < nq; q++) acc += facq[q] * stage[q * ND + lane];
            __syncwarp();
          }
        }
      }
      if constexpr (SINK == 0) {
        if (lane < ND) A.elvecs[k * ND + lane] = acc;
        if (A.npts && lane == 0) A.npts[k] = np_used;
      } else {
        if (active && lane < ND) A.vec[A.el2dof_list[k * ND + lane]] += acc;
        __syncwarp();
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// element vectors, ONE THREAD PER LIST ENTRY (orders 1-2, dense sink): the warp-per-element kernel above keeps ~10 of 32 lanes
// busy on an uncut element (3.7 ms for the 942 504 HASNEG P2 tets of a 4.98 M-tet mesh).  Two passes like the deformed kernels:
// pass 1 integrates the uncut entries of the domain (ND accumulators in registers, rule points as shared-memory broadcasts) and
// records the cut ones (integer slot from an atomic counter; every vector is written at its own entry, so the order of the list
// never reaches a value); pass 2 walks the recorded entries through the sub-simplex rules.
// ---------------------------------------------------------------------------------------------------------------
struct ElvecTArgs {
  ElvecArgs E;
  int32_t *work; int32_t *work_count;
  // optional GridFunction factor of the coefficient (H1 order CFO = 1, 2; orientation free): f(x) = poly(x) * sum_i c[dof_i] phi_i(x)
  const double *cf_vec; const int32_t *cf_el2dof;
};

template <int DIM>
__device__ __forceinline__ double poly_at(const PolyDev2 &f, const double *lam, const double (*p)[DIM]) {
  double x[3] = {0.0, 0.0, 0.0};
#pragma unroll
  for (int d = 0; d < DIM; d++) {
    double v = 0.0;
#pragma unroll
    for (int kk = 0; kk <= DIM; kk++) v += lam[kk] * p[kk][d];
    x[d] = v;
  }
  double fx = 0.0;
  for (int m = 0; m < f.n; m++) fx += f.c[m] * ipow(x[0], f.ex[m][0]) * ipow(x[1], f.ex[m][1]) * ipow(x[2], f.ex[m][2]);
  return fx;
}

// value of the GridFunction factor at barycentrics lam (CFO = 0: no such factor)
template <int DIM, int CFO>
__device__ __forceinline__ double gf_factor(const double *cfl, const double *lam) {
  if constexpr (CFO == 0) return 1.0;
  else {
    double v = 0.0;
#pragma unroll
    for (int i = 0; i < H1<DIM, CFO>::ND; i++) v += cfl[i] * h1_shape<DIM, CFO>(i, lam);
    return v;
  }
}

// test functions of the element vector; ORDER 0: the constant 1, i.e. the element's share of a scalar integral (Integrate)
template <int DIM, int ORDER> struct TestFn {
  static constexpr int ND = H1<DIM, ORDER == 0 ? 1 : ORDER>::ND;
  static __device__ __forceinline__ double value(int i, const double *lam) { return h1_shape<DIM, ORDER == 0 ? 1 : ORDER>(i, lam); }
};
template <int DIM> struct TestFn<DIM, 0> {
  static constexpr int ND = 1;
  static __device__ __forceinline__ double value(int, const double *) { return 1.0; }
};

template <int DIM, int ORDER, bool CUTPASS, int CFO>
__global__ void __launch_bounds__(128) k_elvec_t(ElvecTArgs T, int64_t nlist) {
  constexpr int ND = TestFn<DIM, ORDER>::ND;
  const ElvecArgs &A = T.E;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double4 *srule_vol = reinterpret_cast<double4 *>(smem_raw);
  double4 *srule_if = srule_vol + A.nr_vol;
  CutEntry *lut = reinterpret_cast<CutEntry *>(srule_if + A.nr_if);
  for (int i = threadIdx.x; i < A.nr_vol; i += blockDim.x) srule_vol[i] = A.rule_vol[i];
  for (int i = threadIdx.x; i < A.nr_if; i += blockDim.x) srule_if[i] = A.rule_if[i];
  if (threadIdx.x < 16) lut[threadIdx.x] = A.tab->e[DIM - 2][A.dt][threadIdx.x];
  __syncthreads();
  const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t k = w;
  if constexpr (CUTPASS) {
    if (w >= (int64_t)*T.work_count) return;
    k = T.work[w];
  } else if (w >= nlist) return;
  const int64_t e = A.els ? A.els[k] : k;
  double ls[4], p[DIM + 1][DIM];
  load_el<DIM>(A.M, A.lset, e, ls, p);
  int mask;
  const int dom = sign_test<DIM>(ls, mask);
  double acc[ND];
#pragma unroll
  for (int i = 0; i < ND; i++) acc[i] = 0.0;
  constexpr int NCF = CFO == 0 ? 1 : H1<DIM, CFO == 0 ? 1 : CFO>::ND;
  double cfl[NCF];
  if constexpr (CFO > 0) {
#pragma unroll
    for (int i = 0; i < NCF; i++) cfl[i] = __ldg(T.cf_vec + __ldg(T.cf_el2dof + e * NCF + i));
  }
  int np_used = 0;
  if constexpr (!CUTPASS) {
    if (dom == IF) {
      T.work[atomicAdd(T.work_count, 1)] = (int32_t)k;
      return;
    }
    if (dom == A.dt) {
      np_used = A.nr_vol;
      for (int q = 0; q < A.nr_vol; q++) {
        const double4 r = srule_vol[q];
        double lam[4] = {r.x, r.y, r.z, 0.0};
        double last = 1.0;
#pragma unroll
        for (int kk = 0; kk < DIM; kk++) last -= lam[kk];
        lam[DIM] = last;
        const double fw = r.w * poly_at<DIM>(A.f, lam, p) * gf_factor<DIM, CFO>(cfl, lam);
#pragma unroll
        for (int i = 0; i < ND; i++) acc[i] += fw * TestFn<DIM, ORDER>::value(i, lam);
      }
    }
  } else {
    ElGeom<DIM> G;
    element_geometry<DIM>(p, G);
    double scale = 1.0;
    if (A.dt == IF) {
      double gl[DIM + 1][DIM];
      bary_gradients<DIM>(G, gl);
      double g2 = 0.0, r2 = 0.0;
#pragma unroll
      for (int d = 0; d < DIM; d++) {
        double v = 0.0;
#pragma unroll
        for (int kk = 0; kk <= DIM; kk++) v += ls[kk] * gl[kk][d];
        g2 += v * v;
      }
#pragma unroll
      for (int kk = 0; kk < DIM; kk++) { double v = ls[kk] - ls[DIM]; r2 += v * v; }
      scale = sqrt(g2) / sqrt(r2);
    }
    const CutEntry E = lut[mask];
    const double4 *rule = A.dt == IF ? srule_if : srule_vol;
    const int nr = A.dt == IF ? A.nr_if : A.nr_vol;
    np_used = E.ns * nr;
    for (int s = 0; s < E.ns; s++) {
      double P[4][3];
      for (int v = 0; v < E.npt; v++) coded_point<DIM>(E.v[s][v], ls, P[v]);
      const double sfac = simplex_factor(E.npt, P) * scale;
      for (int q = 0; q < nr; q++) {
        const double4 r = rule[q];
        double xi[3];
        simplex_point(E.npt, P, r, xi);
        double lam[4] = {xi[0], xi[1], xi[2], 0.0};
        double last = 1.0;
#pragma unroll
        for (int kk = 0; kk < DIM; kk++) last -= lam[kk];
        lam[DIM] = last;
        const double fw = (r.w * sfac) * poly_at<DIM>(A.f, lam, p) * gf_factor<DIM, CFO>(cfl, lam);
#pragma unroll
        for (int i = 0; i < ND; i++) acc[i] += fw * TestFn<DIM, ORDER>::value(i, lam);
      }
    }
  }
  double meas = 0.0;
  if (np_used) {
    ElGeom<DIM> G;
    element_geometry<DIM>(p, G);
    meas = fabs(G.det);
  }
  double *o = A.elvecs + k * ND;
#pragma unroll
  for (int i = 0; i < ND; i++) o[i] = meas * acc[i];
  if (A.npts) A.npts[k] = np_used;
}

// fixed-shape sum: block b adds its contiguous chunk (thread-strided, then a tree), one block adds the block sums
__global__ void __launch_bounds__(256) k_sum_chunks(int64_t n, const double *__restrict__ in, double *__restrict__ partial) {
  __shared__ double s[256];
  const int64_t per = (n + gridDim.x - 1) / gridDim.x, lo = (int64_t)blockIdx.x * per, hi = min(n, lo + per);
  double acc = 0.0;
  for (int64_t i = lo + threadIdx.x; i < hi; i += 256) acc += in[i];
  s[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = s[0];
}

__global__ void __launch_bounds__(1024) k_sum_final(int64_t n, const double *__restrict__ partial, double *__restrict__ out) {
  __shared__ double s[1024];
  s[threadIdx.x] = (int64_t)threadIdx.x < n ? partial[threadIdx.x] : 0.0;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = s[0];
}

template <int DIM, int ORDER, int CFO = 0>
static int launch_elvec_t(const ElvecArgs &A, int64_t nlist, cudaStream_t st, const double *cf_vec = nullptr, const int32_t *cf_el2dof = nullptr) {
  if (nlist >= (int64_t)INT32_MAX) { set_error("elvec_h1: more than 2^31 - 1 list entries"); return NGSX_ERR_ARG; }
  int32_t *work = nullptr;
  NGSX_CUDA(cudaMallocAsync(&work, (nlist + 1) * sizeof(int32_t), st));
  ElvecTArgs T{A, work + 1, work, cf_vec, cf_el2dof};
  cudaError_t e0 = cudaMemsetAsync(work, 0, sizeof(int32_t), st);
  const size_t smem = (size_t)(A.nr_vol + A.nr_if) * 32 + 16 * sizeof(CutEntry);
  const unsigned nb = (unsigned)((nlist + 127) / 128);
  k_elvec_t<DIM, ORDER, false, CFO><<<nb, 128, smem, st>>>(T, nlist);
  k_elvec_t<DIM, ORDER, true, CFO><<<nb, 128, smem, st>>>(T, nlist);
  g_launches.fetch_add(2, std::memory_order_relaxed);
  cudaError_t e1 = cudaGetLastError();
  cudaFreeAsync(work, st);
  NGSX_CUDA(e0);
  NGSX_CUDA(e1);
  return NGSX_OK;
}

// slot map: one thread per (list entry, i, j): binary search of the sorted CSR row
__global__ void __launch_bounds__(256) k_slot_map(int64_t n, int nd, const int32_t *__restrict__ table, const int64_t *__restrict__ rowptr,
                                                 const int32_t *__restrict__ colidx, int32_t *__restrict__ slots) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int ne = nd * nd;
  if (t >= n * ne) return;
  int64_t k = t / ne;
  int idx = (int)(t % ne);
  int row = table[k * nd + idx / nd], col = table[k * nd + idx % nd];
  int64_t lo = rowptr[row], hi = rowptr[row + 1] - 1;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (colidx[mid] < col) lo = mid + 1; else hi = mid;
  }
  slots[t] = (colidx[lo] == col) ? (int32_t)lo : -1;
}

template <typename K>
static int configure(K kernel, size_t smem, int sm_count, int64_t nwarp_items, unsigned *grid) {
  NGSX_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, WARPS * 32, smem));
  if (occ < 1) { set_error("kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
  int64_t need = (nwarp_items + WARPS - 1) / WARPS;
  int64_t cap = (int64_t)sm_count * occ;
  *grid = (unsigned)std::max<int64_t>(1, std::min(need, cap));
  return NGSX_OK;
}

static int fill_common(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, const DeviceData **dd,
                       const double4 **rv, int *nv, const double4 **ri, int *ni, int max_order = 2) {
  NGSX_REQUIRE(mesh && (mesh->dim == 2 || mesh->dim == 3), "mesh.dim must be 2 or 3");
  NGSX_REQUIRE(lset_v && mesh->vert && mesh->el2v, "mesh / level-set arrays are NULL");
  NGSX_REQUIRE(dt >= 0 && dt <= 2, "domain_type must be NEG, POS or IF");
  if (fe_order < 1 || fe_order > max_order) {
    set_error(max_order == 3 ? "H1 order outside the CUDA path (supported: 1, 2, 3); no CPU fallback"
                             : "H1 order outside this entry point (supported: 1, 2; order 3: ngsx_elmat_h1 / ngsx_assemble_csr_h1); no CPU fallback");
    return NGSX_ERR_UNSUPPORTED;
  }
  *dd = device_data();
  if (!*dd) return NGSX_ERR_CUDA;
  if (!rule_lookup(*dd, mesh->dim, intorder, rv, nv) || !rule_lookup(*dd, mesh->dim - 1, intorder, ri, ni)) {
    set_error("integration order beyond the tabulated reference rules");
    return NGSX_ERR_UNSUPPORTED;
  }
  return NGSX_OK;
}

static int g_elvec_warp = 0;      // debug: dense element vectors through the warp-per-element kernel k_elvec (A/B and parity)
static int g_order3_fma = 0;      // debug: order 3 through the first-generation FMA kernel k_elmat<DIM,3> (A/B and parity)

template <int DIM, int SINK>
static int launch_elmat3(const ElmatArgs &A, int sm_count, int ncolors, const int64_t *h_color_ptr, int64_t nlist, cudaStream_t st) {
  using S = Smem3<DIM>;
  size_t smem = S::total(A.nr_vol, A.nr_if);
  auto kern = k_elmat3<DIM, SINK>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, W3 * 32, smem));
  if (occ < 1) { set_error("kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
  if constexpr (SINK == 0) {
    if (A.use_fast && A.dt != IF && !A.sublist && !A.d_count && !h_color_ptr && nlist < (int64_t)INT32_MAX) {
      // two passes: table blocks of the uncut entries (light kernel, full occupancy), then the recorded cut entries on the tensor cores
      int32_t *work = nullptr;
      NGSX_CUDA(cudaMallocAsync(&work, (nlist + 1) * sizeof(int32_t), st));
      ElmatArgs B = A;
      B.work = work + 1; B.work_count = work;
      cudaError_t e0 = cudaMemsetAsync(work, 0, sizeof(int32_t), st);
      constexpr int NTRI = H1<DIM, 3>::ND * (H1<DIM, 3>::ND + 1) / 2, NPAIR = DIM * (DIM + 1) / 2;
      const size_t smem1 = (size_t)A.nr_vol * 32 + (size_t)(NPAIR + 1) * (NTRI + H1<DIM, 3>::ND * H1<DIM, 3>::ND) * 8;
      auto k1 = k_elmat3_fast<DIM>;
      int occ1 = 0;
      cudaError_t e1 = cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1);
      if (e1 == cudaSuccess) e1 = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ1, k1, W3F * 32, smem1);
      if (e1 == cudaSuccess && e0 == cudaSuccess && occ1 >= 1) {
        const unsigned g1 = (unsigned)std::max<int64_t>(1, std::min<int64_t>(((nlist + 1) / 2 + W3F - 1) / W3F, (int64_t)sm_count * occ1));
        k1<<<g1, W3F * 32, smem1, st>>>(B, nlist);
        const unsigned g2 = (unsigned)std::max<int64_t>(1, std::min<int64_t>((nlist + W3 - 1) / W3, (int64_t)sm_count * occ));
        kern<<<g2, W3 * 32, smem, st>>>(B, 0, nlist);
        g_launches.fetch_add(2, std::memory_order_relaxed);
        e1 = cudaGetLastError();
      }
      cudaFreeAsync(work, st);
      NGSX_CUDA(e0);
      NGSX_CUDA(e1);
      if (occ1 < 1) { set_error("kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
      return NGSX_OK;
    }
  }
  for (int c = 0; c < ncolors; c++) {
    int64_t first = h_color_ptr ? h_color_ptr[c] : 0;
    int64_t count = h_color_ptr ? h_color_ptr[c + 1] - h_color_ptr[c] : nlist;
    if (count <= 0) continue;
    const int64_t need = ((count + A.group - 1) / A.group + W3 - 1) / W3;
    const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)sm_count * occ));
    kern<<<grid, W3 * 32, smem, st>>>(A, first, count);
    NGSX_LAUNCH_CHECK();
  }
  return NGSX_OK;
}

template <int DIM, int ORDER, int SINK>
static int launch_elmat(const ElmatArgs &A, int sm_count, int ncolors, const int64_t *h_color_ptr, int64_t nlist, cudaStream_t st) {
  if constexpr (ORDER == 3) {
    if (!g_order3_fma) return launch_elmat3<DIM, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
  }
  using S = Smem<DIM, ORDER>;
  size_t smem = S::total(A.nr_vol, A.nr_if);
  auto kern = k_elmat<DIM, ORDER, SINK>;
  for (int c = 0; c < ncolors; c++) {
    int64_t first = h_color_ptr ? h_color_ptr[c] : 0;
    int64_t count = h_color_ptr ? h_color_ptr[c + 1] - h_color_ptr[c] : nlist;
    if (count <= 0) continue;
    unsigned grid;
    if (int rc = configure(kern, smem, sm_count, (count + A.group - 1) / A.group, &grid)) return rc;
    kern<<<grid, WARPS * 32, smem, st>>>(A, first, count);
    NGSX_LAUNCH_CHECK();
  }
  return NGSX_OK;
}

template <int SINK>
static int dispatch_elmat(const ElmatArgs &A, int fe_order, int sm_count, int ncolors, const int64_t *h_color_ptr, int64_t nlist, cudaStream_t st) {
  int key = A.M.dim * 10 + fe_order;
  switch (key) {
    case 21: return launch_elmat<2, 1, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
    case 22: return launch_elmat<2, 2, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
    case 31: return launch_elmat<3, 1, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
    case 32: return launch_elmat<3, 2, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
    case 23: return launch_elmat<2, 3, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
    case 33: return launch_elmat<3, 3, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
  }
  set_error("unsupported (dim, order)");
  return NGSX_ERR_UNSUPPORTED;
}

template <int DIM, int ORDER, int SINK>
static int launch_elvec(const ElvecArgs &A, int sm_count, int ncolors, const int64_t *h_color_ptr, int64_t nlist, cudaStream_t st) {
  constexpr int ND = H1<DIM, ORDER>::ND;
  size_t smem = (size_t)(A.nr_vol + A.nr_if) * 32 + 256 + (size_t)WARPS * (32 * ND + 32) * 8;
  auto kern = k_elvec<DIM, ORDER, SINK>;
  for (int c = 0; c < ncolors; c++) {
    int64_t first = h_color_ptr ? h_color_ptr[c] : 0;
    int64_t count = h_color_ptr ? h_color_ptr[c + 1] - h_color_ptr[c] : nlist;
    if (count <= 0) continue;
    unsigned grid;
    if (int rc = configure(kern, smem, sm_count, (count + A.group - 1) / A.group, &grid)) return rc;
    kern<<<grid, WARPS * 32, smem, st>>>(A, first, count);
    NGSX_LAUNCH_CHECK();
  }
  return NGSX_OK;
}

template <int SINK>
static int dispatch_elvec(const ElvecArgs &A, int fe_order, int sm_count, int ncolors, const int64_t *h_color_ptr, int64_t nlist, cudaStream_t st) {
  int key = A.M.dim * 10 + fe_order;
  switch (key) {
    case 21: return launch_elvec<2, 1, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
    case 22: return launch_elvec<2, 2, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
    case 31: return launch_elvec<3, 1, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
    case 32: return launch_elvec<3, 2, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
    case 23: return launch_elvec<2, 3, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
    case 33: return launch_elvec<3, 3, SINK>(A, sm_count, ncolors, h_color_ptr, nlist, st);
  }
  set_error("unsupported (dim, order)");
  return NGSX_ERR_UNSUPPORTED;
}

// ---------------------------------------------------------------------------------------------------------------
// feeders of the owner-computes row gather (ngsx_rows.cu)
// ---------------------------------------------------------------------------------------------------------------
// one thread per list entry: sign test, then either the geometry record of the constant-rule path or a slot in the
// quadrature work list.  The slot order depends on the atomic counter, the values do not.
// Geometry record of list entry k (ngsx_rows.cuh): per local vertex a the 32 bytes {K[a][0..DIM-1], (pad), gm},
//   K[a][b] = |det J| c_s grad(lam_a).grad(lam_b) (physical gradients: rows of J^-1, grad(lam_DIM) = -sum), gm = |det J| c_m.
// A vertex row of the element matrix needs record a alone, an edge row the records of its two vertices (k_rows).  Entries that
// contribute nothing, or whose matrix comes from the quadrature kernel, carry a NaN-boxed code in the gm slot of every record.
template <int DIM>
__global__ void __launch_bounds__(256) k_geom(ngsx_mesh M, const double *__restrict__ lset, const int32_t *__restrict__ els, int64_t nlist,
                                             int dt, FormDev form, int use_fast, double *__restrict__ G, int32_t *__restrict__ cutlist,
                                             int32_t *__restrict__ cutcount, int64_t cut_cap, unsigned long long *__restrict__ err) {
  // the records of the 32 elements of a warp are one contiguous run of G: staged in shared memory (lane stride 144 bytes:
  // conflict-free 128-bit stores) and written out with coalesced 256-bit stores
  constexpr int NV = DIM + 1, LSTR = NV * 4 + 2;
  __shared__ __align__(32) double sstage[8][32 * LSTR];
  const int lane = threadIdx.x & 31;
  double *stage = sstage[threadIdx.x >> 5];
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t k0 = k - lane;
  if (k0 >= nlist) return;                                               // warp-uniform
  const bool active = k < nlist;
  const int64_t e = active ? (els ? els[k] : k) : 0;
  double ls[4], p[DIM + 1][DIM];
  load_el<DIM>(M, lset, e, ls, p);
  int mask;
  const int dom = active ? sign_test<DIM>(ls, mask) : -2;
  double Kr[DIM + 1][3];
#pragma unroll
  for (int a = 0; a <= DIM; a++) Kr[a][0] = Kr[a][1] = Kr[a][2] = 0.0;
  double gm = __longlong_as_double((long long)GCODE_NAN);                // payload 0: no contribution
  if (dom == -1) {
    if (err) {
      atomicAdd(&err[0], 1ULL);
      atomicMin(reinterpret_cast<long long *>(&err[1]), (long long)e);
    }
  } else if (dom == IF || dom == dt) {
    if (dom != IF && use_fast) {
      ElGeom<DIM> Gm;
      element_geometry<DIM>(p, Gm);
      const double meas = fabs(Gm.det);
      double cm = form.cm, cs = form.cs;
      if (form.pm | form.ps) {
        double h = DIM == 2 ? sqrt(meas) : cbrt(meas);
        cm *= ipow(h, form.pm); cs *= ipow(h, form.ps);
      }
      const double f = meas * cs;
#pragma unroll
      for (int b = 0; b < DIM; b++) {
        double col = 0.0;
#pragma unroll
        for (int a = 0; a < DIM; a++) {
          double s = 0.0;
#pragma unroll
          for (int d = 0; d < DIM; d++) s += Gm.Jinv[a][d] * Gm.Jinv[b][d];
          Kr[a][b] = f * s;
          col += f * s;
        }
        Kr[DIM][b] = -col;
      }
      gm = meas * cm;
    } else {
      const int slot = atomicAdd(cutcount, 1);                           // beyond the capacity: counted (the caller grows the buffer), skipped
      if (slot < cut_cap) {
        cutlist[slot] = (int32_t)k;
        gm = __longlong_as_double((long long)(GCODE_NAN | (unsigned long long)(slot + 1)));
      }
    }
  }
  double2 *mine = reinterpret_cast<double2 *>(stage + lane * LSTR);
#pragma unroll
  for (int a = 0; a <= DIM; a++) { mine[2 * a] = make_double2(Kr[a][0], Kr[a][1]); mine[2 * a + 1] = make_double2(Kr[a][2], gm); }
  __syncwarp();
  const int nrec = (int)min((int64_t)32, nlist - k0) * NV;               // 32-byte records of this warp's run
  double *dst = G + k0 * (NV * 4);
  for (int t = lane; t < nrec; t += 32) {
    const int el = t / NV, a = t - el * NV;
    const double2 *src = reinterpret_cast<const double2 *>(stage + el * LSTR + a * 4);
    const double2 u = src[0], w = src[1];
    asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4 * t), "d"(u.x), "d"(u.y), "d"(w.x), "d"(w.y) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------
// quadrature element matrices, ONE THREAD PER ELEMENT (work list of the row gather: cut elements, Nitsche / IF forms).
// All supported forms are symmetric, so a thread keeps the ND(ND+1)/2 upper-triangle accumulators in registers, walks the
// sub-simplices and their rule points itself (rule and decomposition table are shared-memory broadcasts) and mirrors the
// block on store.  No shared-memory staging, no lane idles on the 14-point rules; the dense nd x nd block of work item s
// goes to elmats[s].
// ---------------------------------------------------------------------------------------------------------------
constexpr int CUTMAT_THREADS = 128;
// dynamic shared memory of the cut-element kernels: rules | decomposition table | per-warp staging of the output blocks
__host__ __device__ inline size_t cutmat_stage_offset(int nr_vol, int nr_if) {
  return ((size_t)(nr_vol + nr_if) * sizeof(double4) + 16 * sizeof(CutEntry) + 15) & ~(size_t)15;
}
// the staged blocks of `nvalid` consecutive work items -> their contiguous run in global memory, coalesced
template <int NE, int ST>
__device__ __forceinline__ void cutmat_flush(const double *stage, double *dst, int nvalid, int lane) {
  __syncwarp();
  const int total = nvalid * NE;
  for (int t = lane; t < total; t += 32) {
    const int el = t / NE, c = t - el * NE;
    dst[t] = stage[el * ST + c];
  }
  __syncwarp();
}

template <int DIM, int ORDER, bool IFACE>
__global__ void __launch_bounds__(CUTMAT_THREADS) k_cutmat(ElmatArgs A, int64_t count) {
  constexpr int ND = H1<DIM, ORDER>::ND, NE = ND * ND, NS = ND * (ND + 1) / 2, NEDGE = ND - (DIM + 1);
  constexpr int NPT = IFACE ? DIM : DIM + 1;       // vertices of a sub-simplex: compile time, so they stay in registers
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double4 *srule_vol = reinterpret_cast<double4 *>(smem_raw);
  double4 *srule_if = srule_vol + A.nr_vol;
  CutEntry *lut = reinterpret_cast<CutEntry *>(srule_if + A.nr_if);
  for (int i = threadIdx.x; i < A.nr_vol; i += blockDim.x) srule_vol[i] = A.rule_vol[i];
  for (int i = threadIdx.x; i < A.nr_if; i += blockDim.x) srule_if[i] = A.rule_if[i];
  if (threadIdx.x < 16) lut[threadIdx.x] = A.tab->e[DIM - 2][A.dt][threadIdx.x];
  __syncthreads();
  if (A.d_count) count = min(count, (int64_t)*A.d_count);      // work list on the device, clamped to the caller's capacity
  // the blocks of the 32 work items of a warp are one contiguous run of elmats: every lane writes its block into the warp's
  // staging area (row stride odd: conflict-free) and the warp streams the run out with coalesced stores.  (Stored from the
  // lanes directly, every store instruction touched 32 sectors in 32 lines and the kernel sat in the LSU queue.)
  constexpr int ST = NE | 1;
  const int lane = threadIdx.x & 31;
  double *stage = reinterpret_cast<double *>(smem_raw + cutmat_stage_offset(A.nr_vol, A.nr_if)) + (size_t)(threadIdx.x >> 5) * 32 * ST;
  for (int64_t s0 = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31); s0 < count; s0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = s0 + lane;
    if (s < count) {
    const int64_t kk = A.sublist ? A.sublist[s] : s;
    const int64_t e = A.els ? A.els[kk] : kk;
    double ls[4], p[DIM + 1][DIM];
    load_el<DIM>(A.M, A.lset, e, ls, p);
    int mask;
    const int dom = sign_test<DIM>(ls, mask);
    const bool cut = dom == IF;
    double acc[NS];
#pragma unroll
    for (int t = 0; t < NS; t++) acc[t] = 0.0;
    if (cut || dom == A.dt) {
      ElGeom<DIM> G;
      element_geometry<DIM>(p, G);
      const double meas = fabs(G.det);
      double cm = A.form.cm, cs = A.form.cs, cn = A.form.cn, cd = A.form.cd;
      if (A.form.pm | A.form.ps | A.form.pn | A.form.pd) {
        double h = DIM == 2 ? sqrt(meas) : cbrt(meas);
        cm *= ipow(h, A.form.pm); cs *= ipow(h, A.form.ps); cn *= ipow(h, A.form.pn); cd *= ipow(h, A.form.pd);
      }
      double gl[DIM + 1][DIM];
      bary_gradients<DIM>(G, gl);
      double nrm[DIM], scale = 1.0;
      {
        double g2 = 0.0, r2 = 0.0;
#pragma unroll
        for (int d = 0; d < DIM; d++) {
          double v = 0.0;
#pragma unroll
          for (int k = 0; k <= DIM; k++) v += ls[k] * gl[k][d];
          nrm[d] = v; g2 += v * v;
        }
#pragma unroll
        for (int k = 0; k < DIM; k++) { double v = ls[k] - ls[DIM]; r2 += v * v; }
        const double inv = 1.0 / sqrt(g2);
#pragma unroll
        for (int d = 0; d < DIM; d++) nrm[d] *= inv;
        if (cut && IFACE) scale = sqrt(g2) / sqrt(r2);                // |F^-T n_ref|
      }
      // normal derivatives of the barycentrics (Nitsche terms)
      double gln[DIM + 1];
#pragma unroll
      for (int k = 0; k <= DIM; k++) {
        double v = 0.0;
#pragma unroll
        for (int d = 0; d < DIM; d++) v += gl[k][d] * nrm[d];
        gln[k] = v;
      }
      CutEntry E;
      if (cut) E = lut[mask]; else { E.ns = 1; E.npt = 0; }
      const double4 *rule = (cut && IFACE) ? srule_if : srule_vol;
      const int nr = (cut && IFACE) ? A.nr_if : A.nr_vol;
      for (int sub = 0; sub < E.ns; sub++) {
        double P[4][3];
        double sfac = 1.0;
        if (cut) {
#pragma unroll
          for (int v = 0; v < NPT; v++) coded_point<DIM>(E.v[sub][v], ls, P[v]);
          sfac = simplex_factor(NPT, P);
        }
        const double f0 = meas * sfac * scale;
        for (int q = 0; q < nr; q++) {
          const double4 r = rule[q];
          double xi[3] = {r.x, r.y, r.z};
          if (cut) simplex_point(NPT, P, r, xi);
          double lam[DIM + 1];
          double last = 1.0;
#pragma unroll
          for (int k = 0; k < DIM; k++) { lam[k] = xi[k]; last -= xi[k]; }
          lam[DIM] = last;
          const double f = f0 * r.w;
          // B rows: value, gradient, normal derivative of every shape function at this point
          double sv[ND], gv[ND][DIM], gn[ND];
#pragma unroll
          for (int i = 0; i <= DIM; i++) {
            sv[i] = lam[i]; gn[i] = gln[i];
#pragma unroll
            for (int d = 0; d < DIM; d++) gv[i][d] = gl[i][d];
          }
          if constexpr (ORDER == 2) {
#pragma unroll
            for (int ed = 0; ed < NEDGE; ed++) {
              int a, b;
              edge_vertices<DIM>(ed, a, b);
              const int i = DIM + 1 + ed;
              sv[i] = -0.5 * lam[a] * lam[b];
              gn[i] = -0.5 * (gln[a] * lam[b] + lam[a] * gln[b]);
#pragma unroll
              for (int d = 0; d < DIM; d++) gv[i][d] = -0.5 * (gl[a][d] * lam[b] + lam[a] * gl[b][d]);
            }
          }
          const double fs = f * cs, fm = f * cm, fn = f * cn, fd = f * cd;
          int t = 0;
#pragma unroll
          for (int i = 0; i < ND; i++) {
            double ag[DIM];
#pragma unroll
            for (int d = 0; d < DIM; d++) ag[d] = fs * gv[i][d];
            // (cm phi_i + cn dn phi_i) phi_j + (cn phi_i + cd dn phi_i) dn phi_j
            const double am = fm * sv[i] + fn * gn[i], an = fn * sv[i] + fd * gn[i];
#pragma unroll
            for (int j = 0; j <= i; j++) {
              double v = fma(am, sv[j], acc[t]);
#pragma unroll
              for (int d = 0; d < DIM; d++) v = fma(ag[d], gv[j][d], v);
              if (cn != 0.0 || cd != 0.0) v = fma(an, gn[j], v);
              acc[t++] = v;
            }
          }
        }
      }
    }
    double *out = stage + lane * ST;
    int t = 0;
#pragma unroll
    for (int i = 0; i < ND; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) {
        out[i * ND + j] = acc[t];
        if (j != i) out[j * ND + i] = acc[t];
        t++;
      }
    }
    cutmat_flush<NE, ST>(stage, A.elmats + s0 * NE, (int)min((int64_t)32, count - s0), lane);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Same matrices from MONOMIAL MOMENTS (forms without normal-derivative terms).  The hierarchical shapes are lam_i and
// -1/2 lam_a lam_b, their barycentric derivatives are constants and -1/2 lam, so every entry of the mass and stiffness
// block is a fixed combination of int lam^alpha over the cut part: per rule point the thread accumulates the moments
// (3D P2: 1 + 4 + 10 + 24 + 21 = 60 accumulators, 76 FP64 instructions) instead of the 55 x 4 entry updates plus the
// shape / gradient evaluation (~ 340), and expands them into the nd x nd block once per element.  Same rule, same points
// and weights; only the summation order differs (1e-15 relative).
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ constexpr int sym2(int a, int b) { return a >= b ? a * (a + 1) / 2 + b : b * (b + 1) / 2 + a; }

template <int DIM, int ORDER, bool IFACE>
__global__ void __launch_bounds__(CUTMAT_THREADS) k_cutmat_mom(ElmatArgs A, int64_t count) {
  constexpr int ND = H1<DIM, ORDER>::ND, NE = ND * ND, NV = DIM + 1, NEDGE = ND - NV;
  constexpr int N2 = NV * (NV + 1) / 2, NT = NEDGE > 0 ? NEDGE : 1, NQ = NEDGE > 0 ? NEDGE * (NEDGE + 1) / 2 : 1;
  constexpr int NPT = IFACE ? DIM : DIM + 1;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double4 *srule_vol = reinterpret_cast<double4 *>(smem_raw);
  double4 *srule_if = srule_vol + A.nr_vol;
  CutEntry *lut = reinterpret_cast<CutEntry *>(srule_if + A.nr_if);
  for (int i = threadIdx.x; i < A.nr_vol; i += blockDim.x) srule_vol[i] = A.rule_vol[i];
  for (int i = threadIdx.x; i < A.nr_if; i += blockDim.x) srule_if[i] = A.rule_if[i];
  if (threadIdx.x < 16) lut[threadIdx.x] = A.tab->e[DIM - 2][A.dt][threadIdx.x];
  __syncthreads();
  if (A.d_count) count = min(count, (int64_t)*A.d_count);      // work list on the device, clamped to the caller's capacity
  // the blocks of the 32 work items of a warp are one contiguous run of elmats: every lane writes its block into the warp's
  // staging area (row stride odd: conflict-free) and the warp streams the run out with coalesced stores.  (Stored from the
  // lanes directly, every store instruction touched 32 sectors in 32 lines and the kernel sat in the LSU queue.)
  constexpr int ST = NE | 1;
  const int lane = threadIdx.x & 31;
  double *stage = reinterpret_cast<double *>(smem_raw + cutmat_stage_offset(A.nr_vol, A.nr_if)) + (size_t)(threadIdx.x >> 5) * 32 * ST;
  for (int64_t s0 = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31); s0 < count; s0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = s0 + lane;
    if (s < count) {
    const int64_t kk = A.sublist ? A.sublist[s] : s;
    const int64_t e = A.els ? A.els[kk] : kk;
    double ls[4], p[DIM + 1][DIM];
    load_el<DIM>(A.M, A.lset, e, ls, p);
    int mask;
    const int dom = sign_test<DIM>(ls, mask);
    const bool cut = dom == IF;
    double m0 = 0.0, m1[NV], m2[N2], mT[NV][NT], mQ[NQ];
#pragma unroll
    for (int a = 0; a < NV; a++) {
      m1[a] = 0.0;
#pragma unroll
      for (int t = 0; t < NT; t++) mT[a][t] = 0.0;
    }
#pragma unroll
    for (int t = 0; t < N2; t++) m2[t] = 0.0;
#pragma unroll
    for (int t = 0; t < NQ; t++) mQ[t] = 0.0;
    double cm = A.form.cm, cs = A.form.cs;
    double gl[DIM + 1][DIM];
#pragma unroll
    for (int k = 0; k <= DIM; k++)
#pragma unroll
      for (int d = 0; d < DIM; d++) gl[k][d] = 0.0;
    if (cut || dom == A.dt) {
      ElGeom<DIM> G;
      element_geometry<DIM>(p, G);
      const double meas = fabs(G.det);
      if (A.form.pm | A.form.ps) {
        double h = DIM == 2 ? sqrt(meas) : cbrt(meas);
        cm *= ipow(h, A.form.pm); cs *= ipow(h, A.form.ps);
      }
      bary_gradients<DIM>(G, gl);
      double scale = 1.0;
      if (cut && IFACE) {                                              // |F^-T n_ref|
        double g2 = 0.0, r2 = 0.0;
#pragma unroll
        for (int d = 0; d < DIM; d++) {
          double v = 0.0;
#pragma unroll
          for (int k = 0; k <= DIM; k++) v += ls[k] * gl[k][d];
          g2 += v * v;
        }
#pragma unroll
        for (int k = 0; k < DIM; k++) { double v = ls[k] - ls[DIM]; r2 += v * v; }
        scale = sqrt(g2) / sqrt(r2);
      }
      CutEntry E;
      if (cut) E = lut[mask]; else { E.ns = 1; E.npt = 0; }
      const double4 *rule = (cut && IFACE) ? srule_if : srule_vol;
      const int nr = (cut && IFACE) ? A.nr_if : A.nr_vol;
      for (int sub = 0; sub < E.ns; sub++) {
        double P[4][3];
        double sfac = 1.0;
        if (cut) {
#pragma unroll
          for (int v = 0; v < NPT; v++) coded_point<DIM>(E.v[sub][v], ls, P[v]);
          sfac = simplex_factor(NPT, P);
        }
        const double f0 = meas * sfac * scale;
        for (int q = 0; q < nr; q++) {
          const double4 r = rule[q];
          double xi[3] = {r.x, r.y, r.z};
          if (cut) simplex_point(NPT, P, r, xi);
          double lam[NV], fl[NV];
          double last = 1.0;
#pragma unroll
          for (int k = 0; k < DIM; k++) { lam[k] = xi[k]; last -= xi[k]; }
          lam[DIM] = last;
          const double f = f0 * r.w;
          m0 += f;
#pragma unroll
          for (int a = 0; a < NV; a++) {
            fl[a] = f * lam[a];
#pragma unroll
            for (int b = 0; b <= a; b++) m2[a * (a + 1) / 2 + b] = fma(fl[a], lam[b], m2[a * (a + 1) / 2 + b]);
          }
          if constexpr (ORDER == 2) {
            double ep[NT];
#pragma unroll
            for (int ed = 0; ed < NEDGE; ed++) {
              int a, b;
              edge_vertices<DIM>(ed, a, b);
              ep[ed] = lam[a] * lam[b];
            }
#pragma unroll
            for (int a = 0; a < NV; a++) {
              m1[a] += fl[a];
#pragma unroll
              for (int ed = 0; ed < NEDGE; ed++) mT[a][ed] = fma(fl[a], ep[ed], mT[a][ed]);
            }
            int t = 0;
#pragma unroll
            for (int ed = 0; ed < NEDGE; ed++) {
              const double fe = f * ep[ed];
#pragma unroll
              for (int e2 = 0; e2 <= ed; e2++) { mQ[t] = fma(fe, ep[e2], mQ[t]); t++; }
            }
          }
        }
      }
    }
    // expand: gram matrix of the barycentric gradients, then the lower triangle (mirrored on store)
    double gm[N2];
#pragma unroll
    for (int a = 0; a < NV; a++)
#pragma unroll
      for (int b = 0; b <= a; b++) {
        double v = 0.0;
#pragma unroll
        for (int d = 0; d < DIM; d++) v = fma(gl[a][d], gl[b][d], v);
        gm[a * (a + 1) / 2 + b] = cs * v;
      }
    double *out = stage + lane * ST;
#pragma unroll
    for (int i = 0; i < NV; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) {
        const double v = fma(cm, m2[sym2(i, j)], gm[sym2(i, j)] * m0);
        out[i * ND + j] = v;
        if (j != i) out[j * ND + i] = v;
      }
    if constexpr (ORDER == 2) {
      int t = 0;
#pragma unroll
      for (int ed = 0; ed < NEDGE; ed++) {
        int a, b;
        edge_vertices<DIM>(ed, a, b);
        const int i = NV + ed;
#pragma unroll
        for (int j = 0; j < NV; j++) {
          const double st = fma(gm[sym2(j, a)], m1[b], gm[sym2(j, b)] * m1[a]);
          const double v = -0.5 * fma(cm, mT[j][ed], st);
          out[i * ND + j] = v;
          out[j * ND + i] = v;
        }
#pragma unroll
        for (int e2 = 0; e2 <= ed; e2++) {
          int c, d;
          edge_vertices<DIM>(e2, c, d);
          double st = gm[sym2(a, c)] * m2[sym2(b, d)];
          st = fma(gm[sym2(a, d)], m2[sym2(b, c)], st);
          st = fma(gm[sym2(b, c)], m2[sym2(a, d)], st);
          st = fma(gm[sym2(b, d)], m2[sym2(a, c)], st);
          const double v = 0.25 * fma(cm, mQ[t], st);
          t++;
          out[i * ND + NV + e2] = v;
          if (e2 != ed) out[(NV + e2) * ND + i] = v;
        }
      }
    }
    }
    cutmat_flush<NE, ST>(stage, A.elmats + s0 * NE, (int)min((int64_t)32, count - s0), lane);
  }
}

static int g_use_fast = 1;

template <int DIM, int ORDER, bool IFACE>
static int launch_cutmat_t(const ElmatArgs &A, int sm_count, int64_t maxcount, cudaStream_t st) {
  constexpr int NE = H1<DIM, ORDER>::ND * H1<DIM, ORDER>::ND;
  size_t smem = cutmat_stage_offset(A.nr_vol, A.nr_if) + (size_t)(CUTMAT_THREADS / 32) * 32 * (NE | 1) * sizeof(double);
  const bool mom = g_use_fast && A.form.cn == 0.0 && A.form.cd == 0.0;
  auto kern = mom ? k_cutmat_mom<DIM, ORDER, IFACE> : k_cutmat<DIM, ORDER, IFACE>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, CUTMAT_THREADS, smem));
  if (occ < 1) { set_error("cut-element kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
  int64_t need = (maxcount + CUTMAT_THREADS - 1) / CUTMAT_THREADS;
  unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)sm_count * occ));
  kern<<<grid, CUTMAT_THREADS, smem, st>>>(A, maxcount);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

static int launch_cutmat(const ElmatArgs &A, int fe_order, int sm_count, int64_t maxcount, cudaStream_t st) {
  switch (A.M.dim * 100 + fe_order * 10 + (A.dt == IF ? 1 : 0)) {
    case 210: return launch_cutmat_t<2, 1, false>(A, sm_count, maxcount, st);
    case 211: return launch_cutmat_t<2, 1, true>(A, sm_count, maxcount, st);
    case 220: return launch_cutmat_t<2, 2, false>(A, sm_count, maxcount, st);
    case 221: return launch_cutmat_t<2, 2, true>(A, sm_count, maxcount, st);
    case 310: return launch_cutmat_t<3, 1, false>(A, sm_count, maxcount, st);
    case 311: return launch_cutmat_t<3, 1, true>(A, sm_count, maxcount, st);
    case 320: return launch_cutmat_t<3, 2, false>(A, sm_count, maxcount, st);
    case 321: return launch_cutmat_t<3, 2, true>(A, sm_count, maxcount, st);
  }
  set_error("unsupported (dim, order)");
  return NGSX_ERR_UNSUPPORTED;
}

// optional per-kernel timing of the row-gather assembly (bench.py roofline): CUDA events on the launch stream
static int g_timing = 0;
constexpr int TIMING_RING = 64;
static cudaEvent_t g_evr[TIMING_RING][4];
static cudaEvent_t *g_ev = g_evr[0];
static int g_ev_calls = 0;      // timed calls since the last ngsx_debug_get_timing
static bool g_ev_ready = false;
// a timing event on the launch stream; under stream capture an EXTERNAL event-record node, so that the events of a replayed
// CUDA graph carry the times of the last replay and cudaEventElapsedTime stays legal
static void record_timing_event(cudaEvent_t e, cudaStream_t st) {
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cs) == cudaSuccess && cs == cudaStreamCaptureStatusActive) cudaEventRecordWithFlags(e, st, cudaEventRecordExternal);
  else cudaEventRecord(e, st);
}
static thread_local cudaEvent_t g_gather_wait = nullptr;      // one-shot: see ngsx_set_gather_wait_event

}  // namespace ngsx

using namespace ngsx;

extern "C" {

// test hook: 0 disables the constant-rule path so both paths can be compared on the GPU
void ngsx_debug_set_fast_path(int on) { g_use_fast = on; }
void ngsx_debug_set_order3_fma(int on) { g_order3_fma = on; }
void ngsx_debug_set_elvec_warp(int on) { g_elvec_warp = on; }

// measurement hook: when on, ngsx_assemble_csr_rows_h1 records events around its kernels (a ring of 64 event sets, so the
// caller need not synchronise every step); ngsx_debug_get_timing waits for the last call and returns the mean over the calls
// since the previous read (at most the last 64) of {geometry factors (+ tables), cut-element matrices, row gather} in ms
void ngsx_debug_set_timing(int on) {
  g_timing = on;
  g_ev_calls = 0;
  if (on && !g_ev_ready) {
    for (auto &set : g_evr)
      for (auto &e : set) cudaEventCreate(&e);
    g_ev_ready = true;
  }
}
int ngsx_debug_get_timing(double *h_ms) {
  NGSX_REQUIRE(h_ms && g_ev_ready, "debug_get_timing: timing was never enabled");
  NGSX_REQUIRE(g_ev_calls > 0, "debug_get_timing: no timed call since the last read");
  NGSX_CUDA(cudaEventSynchronize(g_ev[3]));
  const int n = std::min(g_ev_calls, TIMING_RING);
  for (int k = 0; k < 3; k++) h_ms[k] = 0.0;
  for (int c = 0; c < n; c++) {
    cudaEvent_t *set = g_evr[((g_ev_calls - 1 - c) % TIMING_RING + TIMING_RING) % TIMING_RING];
    for (int k = 0; k < 3; k++) {
      float ms = 0.f;
      NGSX_CUDA(cudaEventElapsedTime(&ms, set[k], set[k + 1]));
      h_ms[k] += ms / n;
    }
  }
  g_ev_calls = 0;
  return NGSX_OK;
}

static void fill_form(const ngsx_form_h1 *f, FormDev &d) {
  d.cm = f->c_mass; d.cs = f->c_stiff; d.cn = f->c_nitsche; d.cd = f->c_dndn;
  d.pm = f->hpow_mass; d.ps = f->hpow_stiff; d.pn = f->hpow_nitsche; d.pd = f->hpow_dndn;
}

int ngsx_elmat_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, const ngsx_form_h1 *h_form,
                  int64_t nlist, const int32_t *els, double *elmats, int32_t *npts, void *stream) {
  const DeviceData *dd; const double4 *rv, *ri; int nv, ni;
  NGSX_REQUIRE(h_form && (nlist == 0 || elmats) && nlist >= 0, "elmat_h1: bad arguments");
  if (int rc = fill_common(mesh, lset_v, fe_order, dt, intorder, &dd, &rv, &nv, &ri, &ni, 3)) return rc;
  if (nlist == 0) return NGSX_OK;
  ElmatArgs A{};
  A.M = *mesh; A.lset = lset_v; A.tab = dd->cut_table; A.dt = dt;
  A.rule_vol = rv; A.nr_vol = nv; A.rule_if = ri; A.nr_if = ni;
  fill_form(h_form, A.form);
  A.els = els; A.group = 1; A.use_fast = g_use_fast && h_form->c_nitsche == 0.0 && h_form->c_dndn == 0.0;
  A.elmats = elmats; A.npts = npts;
  if (take_packed_output()) { set_error("elmat_h1: packed output is not served by this entry point (measured for order 3: the scattered 160-byte row stores cost the producer more than the packed gather saves)"); return NGSX_ERR_UNSUPPORTED; }
  return dispatch_elmat<0>(A, fe_order, dd->sm_count, 1, nullptr, nlist, as_stream(stream));
}

int ngsx_elmat_h1_cf(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, const ngsx_form_h1 *h_form, int cf_order,
                     const int32_t *cf_el2dof, const double *cf_vec, int64_t nlist, const int32_t *els, double *elmats, int32_t *npts, void *stream) {
  const DeviceData *dd; const double4 *rv, *ri; int nv, ni;
  NGSX_REQUIRE(h_form && (nlist == 0 || elmats) && nlist >= 0 && cf_el2dof && cf_vec, "elmat_h1_cf: bad arguments");
  if (int rc = fill_common(mesh, lset_v, fe_order, dt, intorder, &dd, &rv, &nv, &ri, &ni, 2)) return rc;
  if (cf_order < 1 || cf_order > 2) { set_error("elmat_h1_cf: the coefficient GridFunction must be H1 of order 1 or 2"); return NGSX_ERR_UNSUPPORTED; }
  if (nlist == 0) return NGSX_OK;
  ElmatArgs A{};
  A.M = *mesh; A.lset = lset_v; A.tab = dd->cut_table; A.dt = dt;
  A.rule_vol = rv; A.nr_vol = nv; A.rule_if = ri; A.nr_if = ni;
  fill_form(h_form, A.form);
  A.els = els; A.group = 1; A.use_fast = 0;
  A.elmats = elmats; A.npts = npts;
  A.cf_vec = cf_vec; A.cf_el2dof = cf_el2dof; A.cf_order = cf_order;
  return dispatch_elmat<0>(A, fe_order, dd->sm_count, 1, nullptr, nlist, as_stream(stream));
}

int ngsx_assemble_csr_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder,
                         const ngsx_form_h1 *h_form, int64_t nlist, const int32_t *els, const int32_t *slots,
                         int32_t ncolors, const int64_t *h_color_ptr, int32_t group, double *vals, void *stream) {
  const DeviceData *dd; const double4 *rv, *ri; int nv, ni;
  NGSX_REQUIRE(h_form && nlist >= 0 && ncolors >= 0 && group >= 1 && h_color_ptr, "assemble_csr_h1: bad arguments");
  NGSX_REQUIRE(nlist == 0 || (slots && vals), "assemble_csr_h1: slots / vals are NULL");
  if (int rc = fill_common(mesh, lset_v, fe_order, dt, intorder, &dd, &rv, &nv, &ri, &ni, 3)) return rc;
  if (nlist == 0) return NGSX_OK;
  ElmatArgs A{};
  A.M = *mesh; A.lset = lset_v; A.tab = dd->cut_table; A.dt = dt;
  A.rule_vol = rv; A.nr_vol = nv; A.rule_if = ri; A.nr_if = ni;
  fill_form(h_form, A.form);
  A.els = els; A.group = group; A.use_fast = g_use_fast && h_form->c_nitsche == 0.0 && h_form->c_dndn == 0.0;
  A.slots = slots; A.vals = vals;
  return dispatch_elmat<1>(A, fe_order, dd->sm_count, ncolors, h_color_ptr, nlist, as_stream(stream));
}

void ngsx_set_gather_wait_event(void *event) { g_gather_wait = reinterpret_cast<cudaEvent_t>(event); }

int ngsx_assemble_csr_rows_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder,
                              const ngsx_form_h1 *h_form, int64_t nlist, const int32_t *els, ngsx_rowplan *plan,
                              int accumulate, double *vals, int64_t *err, void *stream) {
  cudaEvent_t gather_wait = g_gather_wait;
  g_gather_wait = nullptr;
  const DeviceData *dd; const double4 *rv, *ri; int nv, ni;
  NGSX_REQUIRE(h_form && plan && nlist >= 0 && nlist == plan->nlist, "assemble_csr_rows_h1: bad arguments (nlist must match the plan)");
  NGSX_REQUIRE(plan->nrows == 0 || vals, "assemble_csr_rows_h1: vals is NULL");
  if (int rc = fill_common(mesh, lset_v, fe_order, dt, intorder, &dd, &rv, &nv, &ri, &ni)) return rc;
  const int dim = mesh->dim;
  const int nd = dim == 2 ? (fe_order == 1 ? 3 : 6) : (fe_order == 1 ? 4 : 10);
  NGSX_REQUIRE(nd == plan->nd, "assemble_csr_rows_h1: the plan was built for another element type");
  cudaStream_t st = as_stream(stream);
  if (err) { if (int rc = reset_err(err, st)) return rc; }
  if (nlist == 0 || plan->nrows == 0) return NGSX_OK;
  if (int rc = rows_scratch(plan, dim)) return rc;
  FormDev form;
  fill_form(h_form, form);
  // the closed-form rows of the gather need a permutation-symmetric reference rule (the tabulated NGSolve simplex rules of the
  // orders P1 / P2 forms use are); otherwise, and for forms with normal derivatives / interface integrals, every contributing
  // element goes through the quadrature kernel
  RuleMom mom{};
  const bool sym_rule = rule_moments(dim, intorder, &mom);
  const int use_fast = g_use_fast && sym_rule && h_form->c_nitsche == 0.0 && h_form->c_dndn == 0.0 && dt != IF;
  // 1. geometry records + quadrature work list, 2. cut-element matrices, 3. row gather
  if (g_timing) { g_ev = g_evr[g_ev_calls % TIMING_RING]; g_ev_calls++; record_timing_event(g_ev[0], st); }
  auto *e = reinterpret_cast<unsigned long long *>(err);
  const unsigned nb = (unsigned)((nlist + 255) / 256);
  auto geom = [&]() -> int {
    NGSX_CUDA(cudaMemsetAsync(plan->cutcount, 0, sizeof(int32_t), st));
    if (dim == 2) k_geom<2><<<nb, 256, 0, st>>>(*mesh, lset_v, els, nlist, dt, form, use_fast, plan->G, plan->cutlist, plan->cutcount, plan->cut_cap, e);
    else k_geom<3><<<nb, 256, 0, st>>>(*mesh, lset_v, els, nlist, dt, form, use_fast, plan->G, plan->cutlist, plan->cutcount, plan->cut_cap, e);
    NGSX_LAUNCH_CHECK();
    return NGSX_OK;
  };
  if (plan->cut_cap == 0 || (!use_fast && plan->cut_cap < nlist)) {
    // First numeric call on this plan: size the cut-element buffer from a counting pass (synchronises once; not capturable).
    // Later calls reuse it; a level set that needs more reports through ngsx_rowplan_cut_usage (count > capacity).
    int64_t want = nlist;
    if (use_fast) {
      if (int rc = geom()) return rc;
      int32_t c = 0;
      NGSX_CUDA(cudaMemcpyAsync(&c, plan->cutcount, sizeof c, cudaMemcpyDeviceToHost, st));
      NGSX_CUDA(cudaStreamSynchronize(st));
      want = (int64_t)c + c / 4 + 1024;
      if (err) { if (int rc = reset_err(err, st)) return rc; }
    }
    if (int rc = rows_reserve_cut(plan, want)) return rc;
  }
  if (int rc = geom()) return rc;
  if (g_timing) record_timing_event(g_ev[1], st);
  ElmatArgs A{};
  A.M = *mesh; A.lset = lset_v; A.tab = dd->cut_table; A.dt = dt;
  A.rule_vol = rv; A.nr_vol = nv; A.rule_if = ri; A.nr_if = ni;
  A.form = form;
  A.els = els; A.sublist = plan->cutlist; A.d_count = plan->cutcount; A.group = 1; A.use_fast = 0;
  A.elmats = plan->cutmats; A.npts = nullptr;
  // the work-list length lives on the device: launch the persistent grid for the worst case, idle CTAs leave at once
  if (int rc = launch_cutmat(A, fe_order, dd->sm_count, plan->cut_cap, st)) return rc;
  if (gather_wait) NGSX_CUDA(cudaStreamWaitEvent(st, gather_wait, 0));      // the gather starts once the caller's side work is done
  if (g_timing) record_timing_event(g_ev[2], st);
  const int rc = launch_rows(plan, dim, mom, accumulate, vals, dd->sm_count, st);
  if (g_timing) record_timing_event(g_ev[3], st);
  return rc;
}

static int fill_poly(const ngsx_poly *f, PolyDev2 &d) {
  NGSX_REQUIRE(f && f->nterms >= 0 && f->nterms <= NGSX_POLY_MAXTERMS, "polynomial: bad number of terms");
  d.n = f->nterms;
  for (int m = 0; m < d.n; m++) { d.c[m] = f->coef[m]; for (int k = 0; k < 3; k++) d.ex[m][k] = f->ex[m][k]; }
  return NGSX_OK;
}

int ngsx_elvec_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, const ngsx_poly *h_f,
                  int64_t nlist, const int32_t *els, double *elvecs, int32_t *npts, void *stream) {
  const DeviceData *dd; const double4 *rv, *ri; int nv, ni;
  NGSX_REQUIRE((nlist == 0 || elvecs) && nlist >= 0, "elvec_h1: bad arguments");
  if (int rc = fill_common(mesh, lset_v, fe_order, dt, intorder, &dd, &rv, &nv, &ri, &ni, 3)) return rc;
  ElvecArgs A{};
  if (int rc = fill_poly(h_f, A.f)) return rc;
  if (nlist == 0) return NGSX_OK;
  A.M = *mesh; A.lset = lset_v; A.tab = dd->cut_table; A.dt = dt;
  A.rule_vol = rv; A.nr_vol = nv; A.rule_if = ri; A.nr_if = ni;
  A.els = els; A.group = 1; A.elvecs = elvecs; A.npts = npts;
  if (!g_elvec_warp)                                       // orders 1-2: thread per entry, two passes (order 3 stays on the warp kernel)
    switch (mesh->dim * 10 + fe_order) {
      case 21: return launch_elvec_t<2, 1>(A, nlist, as_stream(stream));
      case 22: return launch_elvec_t<2, 2>(A, nlist, as_stream(stream));
      case 31: return launch_elvec_t<3, 1>(A, nlist, as_stream(stream));
      case 32: return launch_elvec_t<3, 2>(A, nlist, as_stream(stream));
    }
  return dispatch_elvec<0>(A, fe_order, dd->sm_count, 1, nullptr, nlist, as_stream(stream));
}

int ngsx_elvec_h1_cf(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, const ngsx_poly *h_f, int cf_order,
                     const int32_t *cf_el2dof, const double *cf_vec, int64_t nlist, const int32_t *els, double *elvecs, int32_t *npts, void *stream) {
  const DeviceData *dd; const double4 *rv, *ri; int nv, ni;
  NGSX_REQUIRE((nlist == 0 || elvecs) && nlist >= 0 && cf_el2dof && cf_vec, "elvec_h1_cf: bad arguments");
  if (int rc = fill_common(mesh, lset_v, fe_order, dt, intorder, &dd, &rv, &nv, &ri, &ni, 2)) return rc;
  if (cf_order < 1 || cf_order > 2) { set_error("elvec_h1_cf: the coefficient GridFunction must be H1 of order 1 or 2"); return NGSX_ERR_UNSUPPORTED; }
  ElvecArgs A{};
  if (int rc = fill_poly(h_f, A.f)) return rc;
  if (nlist == 0) return NGSX_OK;
  A.M = *mesh; A.lset = lset_v; A.tab = dd->cut_table; A.dt = dt;
  A.rule_vol = rv; A.nr_vol = nv; A.rule_if = ri; A.nr_if = ni;
  A.els = els; A.group = 1; A.elvecs = elvecs; A.npts = npts;
  cudaStream_t st = as_stream(stream);
  switch (mesh->dim * 100 + fe_order * 10 + cf_order) {
    case 211: return launch_elvec_t<2, 1, 1>(A, nlist, st, cf_vec, cf_el2dof);
    case 212: return launch_elvec_t<2, 1, 2>(A, nlist, st, cf_vec, cf_el2dof);
    case 221: return launch_elvec_t<2, 2, 1>(A, nlist, st, cf_vec, cf_el2dof);
    case 222: return launch_elvec_t<2, 2, 2>(A, nlist, st, cf_vec, cf_el2dof);
    case 311: return launch_elvec_t<3, 1, 1>(A, nlist, st, cf_vec, cf_el2dof);
    case 312: return launch_elvec_t<3, 1, 2>(A, nlist, st, cf_vec, cf_el2dof);
    case 321: return launch_elvec_t<3, 2, 1>(A, nlist, st, cf_vec, cf_el2dof);
    case 322: return launch_elvec_t<3, 2, 2>(A, nlist, st, cf_vec, cf_el2dof);
  }
  return NGSX_ERR_UNSUPPORTED;
}

int ngsx_integrate_cf(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, const ngsx_poly *h_f, int cf_order, const int32_t *cf_el2dof,
                      const double *cf_vec, double *out, void *stream) {
  const DeviceData *dd; const double4 *rv, *ri; int nv, ni;
  NGSX_REQUIRE(out && cf_el2dof && cf_vec, "integrate_cf: bad arguments");
  if (int rc = fill_common(mesh, lset_v, 1, dt, order, &dd, &rv, &nv, &ri, &ni, 2)) return rc;
  if (cf_order < 1 || cf_order > 2) { set_error("integrate_cf: the integrand GridFunction must be H1 of order 1 or 2"); return NGSX_ERR_UNSUPPORTED; }
  ElvecArgs A{};
  if (int rc = fill_poly(h_f, A.f)) return rc;
  cudaStream_t st = as_stream(stream);
  const int64_t ne = mesh->ne;
  if (ne == 0) { NGSX_CUDA(cudaMemsetAsync(out, 0, sizeof(double), st)); return NGSX_OK; }
  // per-element shares (every entry written once, at its own place), then a reduction of fixed shape: bitwise reproducible
  const int64_t nmid = std::min<int64_t>((ne + 255) / 256, 1024);
  double *vals = nullptr;
  NGSX_CUDA(cudaMallocAsync(&vals, (size_t)(ne + nmid) * sizeof(double), st));
  A.M = *mesh; A.lset = lset_v; A.tab = dd->cut_table; A.dt = dt;
  A.rule_vol = rv; A.nr_vol = nv; A.rule_if = ri; A.nr_if = ni;
  A.els = nullptr; A.group = 1; A.elvecs = vals; A.npts = nullptr;
  int rc = NGSX_ERR_UNSUPPORTED;
  switch (mesh->dim * 10 + cf_order) {
    case 21: rc = launch_elvec_t<2, 0, 1>(A, ne, st, cf_vec, cf_el2dof); break;
    case 22: rc = launch_elvec_t<2, 0, 2>(A, ne, st, cf_vec, cf_el2dof); break;
    case 31: rc = launch_elvec_t<3, 0, 1>(A, ne, st, cf_vec, cf_el2dof); break;
    case 32: rc = launch_elvec_t<3, 0, 2>(A, ne, st, cf_vec, cf_el2dof); break;
  }
  cudaError_t e1 = cudaSuccess;
  if (rc == NGSX_OK) {
    k_sum_chunks<<<(unsigned)nmid, 256, 0, st>>>(ne, vals, vals + ne);
    k_sum_final<<<1, 1024, 0, st>>>(nmid, vals + ne, out);
    g_launches.fetch_add(2, std::memory_order_relaxed);
    e1 = cudaGetLastError();
  }
  cudaFreeAsync(vals, st);
  if (rc) return rc;
  NGSX_CUDA(e1);
  return NGSX_OK;
}

int ngsx_assemble_vec_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder,
                         const ngsx_poly *h_f, int64_t nlist, const int32_t *els, const int32_t *el2dof_list,
                         int32_t ncolors, const int64_t *h_color_ptr, int32_t group, double *vec, void *stream) {
  const DeviceData *dd; const double4 *rv, *ri; int nv, ni;
  NGSX_REQUIRE(nlist >= 0 && ncolors >= 0 && group >= 1 && h_color_ptr, "assemble_vec_h1: bad arguments");
  NGSX_REQUIRE(nlist == 0 || (el2dof_list && vec), "assemble_vec_h1: el2dof_list / vec are NULL");
  if (int rc = fill_common(mesh, lset_v, fe_order, dt, intorder, &dd, &rv, &nv, &ri, &ni, 3)) return rc;
  ElvecArgs A{};
  if (int rc = fill_poly(h_f, A.f)) return rc;
  if (nlist == 0) return NGSX_OK;
  A.M = *mesh; A.lset = lset_v; A.tab = dd->cut_table; A.dt = dt;
  A.rule_vol = rv; A.nr_vol = nv; A.rule_if = ri; A.nr_if = ni;
  A.els = els; A.group = group; A.el2dof_list = el2dof_list; A.vec = vec;
  return dispatch_elvec<1>(A, fe_order, dd->sm_count, ncolors, h_color_ptr, nlist, as_stream(stream));
}

int ngsx_slot_map(int64_t n, int32_t nd, const int32_t *table, const int64_t *rowptr, const int32_t *colidx,
                  int32_t *slots, void *stream) {
  NGSX_REQUIRE(n >= 0 && nd > 0 && (n == 0 || (table && rowptr && colidx && slots)), "slot_map: bad arguments");
  if (n == 0) return NGSX_OK;
  int64_t tot = n * nd * nd;
  k_slot_map<<<(unsigned)((tot + 255) / 256), 256, 0, as_stream(stream)>>>(n, nd, table, rowptr, colidx, slots);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

// HOST: greedy colouring of groups (see header).  Colours are bits in per-dof masks, 64 at a time.
int ngsx_color_groups(int64_t n, int32_t nd, const int32_t *h_table, int64_t ndof, int32_t group,
                      int64_t *h_perm, int64_t *h_color_ptr, int32_t cap_colors, int32_t *h_ncolors) {
  NGSX_REQUIRE(n >= 0 && nd > 0 && group >= 1 && ndof >= 0 && h_ncolors && (n == 0 || (h_table && h_perm && h_color_ptr)),
               "color_groups: bad arguments");
  *h_ncolors = 0;
  if (n == 0) { if (h_color_ptr) h_color_ptr[0] = 0; return NGSX_OK; }
  const int64_t ng = (n + group - 1) / group;
  std::vector<int32_t> color(ng, -1);
  std::vector<uint64_t> mask(ndof);
  int64_t left = ng;
  int base = 0, maxcol = -1;
  while (left > 0) {                                  // sweep: 64 colours per pass over the still-uncoloured groups
    std::fill(mask.begin(), mask.end(), 0ULL);
    for (int64_t g = 0; g < ng; g++) {
      if (color[g] >= 0) continue;
      const int64_t kb = g * group, ke = std::min<int64_t>(kb + group, n);
      uint64_t used = 0;
      for (int64_t t = kb * nd; t < ke * nd; t++) used |= mask[h_table[t]];
      if (~used == 0ULL) continue;                    // all 64 colours of this sweep taken: next sweep
      int c = __builtin_ctzll(~used);
      color[g] = base + c;
      maxcol = std::max(maxcol, base + c);
      left--;
      const uint64_t bit = 1ULL << c;
      for (int64_t t = kb * nd; t < ke * nd; t++) mask[h_table[t]] |= bit;
    }
    base += 64;
  }
  const int nc = maxcol + 1;
  if (nc > cap_colors) { set_error("color_groups: more colours than the caller's capacity"); return NGSX_ERR_CAPACITY; }
  std::vector<int64_t> cnt(nc + 1, 0);
  for (int64_t g = 0; g < ng; g++) cnt[color[g] + 1] += std::min<int64_t>(group, n - g * group);
  for (int c = 0; c < nc; c++) cnt[c + 1] += cnt[c];
  for (int c = 0; c <= nc; c++) h_color_ptr[c] = cnt[c];
  std::vector<int64_t> pos(cnt.begin(), cnt.end() - 1);
  // within a colour, full groups first so that every warp's [kb, kb+group) window is one original group
  for (int pass = 0; pass < 2; pass++)
    for (int64_t g = 0; g < ng; g++) {
      const int64_t kb = g * group, ke = std::min<int64_t>(kb + group, n);
      const bool full = (ke - kb) == group;
      if (full != (pass == 0)) continue;
      int64_t &p = pos[color[g]];
      for (int64_t k = kb; k < ke; k++) h_perm[p++] = k;
    }
  *h_ncolors = nc;
  return NGSX_OK;
}

}  // extern "C"
