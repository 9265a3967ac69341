// ngsx_deformed.cu -- SURVEY section 8 row f-4: cut element matrices and Integrate on DEFORMED (isoparametric) elements.
//
// Reference behaviour: mesh.SetDeformation(gf) makes every ElementTransformation the base (affine) map plus a vector H1 field,
// x(xi) = sum_i c_i phi_i(xi)  (NGS/comp/meshaccess.cpp ALE_ElementTransformation; per element the same arithmetic as
// FE_ElementTransformation::CalcJacobian, NGS/fem/elementtransformation.cpp: dxdxi = pointmat * dshape).  The straight-cut rule
// stays a rule on the REFERENCE element (the P1 level set lives there, cutint/straightcutrule.cpp:580-661); what changes is
//   * the measure |det F(xi_q)| and the mapped gradients F(xi_q)^-T grad_ref phi_i are per point
//     (xfem/symboliccutbfi.cpp:142 trafo(*ir), :219-227),
//   * interface weights carry |F(xi_q)^-T n_ref| per point (cutint/straightcutrule.cpp:558-576),
//   * the level-set normal grad(lset)/|grad(lset)| is per point,
//   * specialcf.mesh_size = |det F|^(1/dim) (NGS/fem/python_fem.cpp:564-593) is per point as well: MeshSizeCF is not flagged
//     elementwise constant, so a coefficient tree containing it takes the cf->Evaluate(mir, val) branch of
//     xfem/symboliccutbfi.cpp:205-214 (pinned against the compiled reference: oracle/ref_driver.cpp ElmatForm).
// So there is no constant-rule shortcut: every listed element of the domain takes the quadrature path.
//
// Execution model: persistent CTAs of 8 warps, ONE WARP PER ELEMENT.  The warp parks the element's geometry coefficients
// (vertex coordinates + deformation, NDG x DIM doubles) in its shared-memory slice; lane q evaluates rule point q of the
// current round of 16 (Jacobian, inverse, shapes, mapped gradients, normal derivative) into the warp's stage, then every lane
// accumulates its 1..4 matrix entries over the staged points -- the layout of k_elmat (ngsx_assemble.cu) with a per-point
// geometry.  FP64-pipe bound (about 2x the instructions of the affine quadrature path per point); the blocks leave as one
// coalesced store per element and reach the CSR values through the owner-computes gather (ngsx_assemble_csr_rows_mats).
#include <algorithm>
#include "ngsx_host.cuh"
#include "ngsx_iso.cuh"

namespace ngsx {

constexpr int DW = 8;             // warps per CTA
constexpr int DQ = 16;            // rule points staged per round

struct DefoDev { const int32_t *el2dof; const double *coef; };
struct FormDevD { double c[4]; int p[4]; };      // mass, stiffness, Nitsche pair, dn-dn
struct PolyDevD { int n; double c[NGSX_POLY_MAXTERMS]; int ex[NGSX_POLY_MAXTERMS][3]; };

struct DefArgs {
  ngsx_mesh M;
  DefoDev D;
  const double *lset;
  const CutTable *tab;
  int dt;
  const double4 *rule_vol; int nr_vol;
  const double4 *rule_if;  int nr_if;
  FormDevD form;
  PolyDevD f;
  const int32_t *els;
  double *elmats; int32_t *npts;       // element matrices
  const int32_t *rowslot;              // packed output (ngsx_set_packed_output): row i of block k -> elmats + rowslot[k nd + i] * ndp
  double *partial;                     // Integrate: one value per element
  unsigned long long *err;
  int32_t *worklist; int32_t *wcount;  // thread-per-element passes: list positions of the cut elements, their number
};

// cooperative load of one element: level-set values (all lanes), geometry coefficients into the warp's shared slice
template <int DIM, int DORD>
__device__ __forceinline__ void load_iso(const DefArgs &A, int64_t e, int lane, double *ls, double *gc) {
  constexpr int NDG = H1<DIM, DORD>::ND;
  const int32_t *ev = A.M.el2v + e * (DIM + 1);
#pragma unroll
  for (int k = 0; k <= DIM; k++) ls[k] = __ldg(A.lset + __ldg(ev + k));
  __syncwarp();
  for (int t = lane; t < NDG * DIM; t += 32) {
    const int i = t / DIM, d = t % DIM;
    double c = __ldg(A.D.coef + (int64_t)__ldg(A.D.el2dof + e * NDG + i) * DIM + d);
    if (i <= DIM) c += __ldg(A.M.vert + (int64_t)__ldg(ev + i) * DIM + d);
    gc[t] = c;
  }
  __syncwarp();
}

template <int DIM, int ORDER, int DORD>
__global__ void __launch_bounds__(DW * 32) k_elmat_def(DefArgs A, int64_t count) {
  constexpr int ND = H1<DIM, ORDER>::ND, NE = ND * ND, NDG = H1<DIM, DORD>::ND, T = (NE + 31) / 32;
  constexpr size_t PER_WARP = (size_t)DQ * ND * 32 + (size_t)DQ * ND * 8 + 2 * DQ * 8 + NDG * DIM * 8;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char *wbase = smem_raw + warp * PER_WARP;
  double4 *stage = reinterpret_cast<double4 *>(wbase);
  double *dnq = reinterpret_cast<double *>(wbase + (size_t)DQ * ND * 32);
  double *facq = dnq + DQ * ND;
  double *hq = facq + DQ;                                               // mesh_size at the staged points
  double *gc = hq + DQ;

  int ei[T], ej[T];
#pragma unroll
  for (int t = 0; t < T; t++) {
    int idx = lane + 32 * t;
    ei[t] = idx < NE ? idx / ND : 0;
    ej[t] = idx < NE ? idx % ND : 0;
  }
  const bool hp = (A.form.p[0] | A.form.p[1] | A.form.p[2] | A.form.p[3]) != 0;
  const bool nterm = A.form.c[2] != 0.0 || A.form.c[3] != 0.0;

  const int64_t nwarps = (int64_t)gridDim.x * DW;
  for (int64_t k = (int64_t)blockIdx.x * DW + warp; k < count; k += nwarps) {
    const int64_t e = A.els ? A.els[k] : k;
    double ls[4];
    load_iso<DIM, DORD>(A, e, lane, ls, gc);
    int mask;
    const int dom = sign_test<DIM>(ls, mask);
    const bool cut = dom == IF;
    if (dom < 0 && A.err && lane == 0) {
      atomicAdd(&A.err[0], 1ULL);
      atomicMin(reinterpret_cast<long long *>(&A.err[1]), (long long)e);
    }
    double acc[T];
#pragma unroll
    for (int t = 0; t < T; t++) acc[t] = 0.0;
    int np_used = 0;
    if (cut || dom == A.dt) {
      const double cm0 = A.form.c[0], cs0 = A.form.c[1], cn0 = A.form.c[2], cd0 = A.form.c[3];
      double r2 = 0.0;
#pragma unroll
      for (int kk = 0; kk < DIM; kk++) { double v = ls[kk] - ls[DIM]; r2 += v * v; }
      const bool iface = cut && A.dt == IF;
      CutEntry E;
      if (cut) E = A.tab->e[DIM - 2][A.dt][mask]; else { E.ns = 1; E.npt = 0; }
      const double4 *rule = iface ? A.rule_if : A.rule_vol;
      const int nr = iface ? A.nr_if : A.nr_vol;
      np_used = E.ns * nr;
      for (int s = 0; s < E.ns; s++) {
        double P[4][3];
        double sfac = 1.0;
        if (cut) {
          for (int v = 0; v < E.npt; v++) coded_point<DIM>(E.v[s][v], ls, P[v]);
          sfac = simplex_factor(E.npt, P);
        }
        for (int q0 = 0; q0 < nr; q0 += DQ) {
          const int nq = min(DQ, nr - q0);
          double adet = 0.0;
          if (lane < nq) {
            double4 r = rule[q0 + lane];
            double xi[3] = {r.x, r.y, r.z};
            if (cut) simplex_point(E.npt, P, r, xi);
            double lam[4] = {xi[0], xi[1], xi[2], 0.0};
            double last = 1.0;
#pragma unroll
            for (int kk = 0; kk < DIM; kk++) last -= lam[kk];
            lam[DIM] = last;
            double x[DIM], gl[DIM + 1][DIM], det;
            iso_point<DIM, DORD>(GcWarp{gc}, lam, x, gl, det);
            adet = fabs(det);
            double nrm[DIM], g2 = 0.0;
#pragma unroll
            for (int d = 0; d < DIM; d++) {
              double v = 0.0;
#pragma unroll
              for (int kk = 0; kk <= DIM; kk++) v += ls[kk] * gl[kk][d];
              nrm[d] = v; g2 += v * v;
            }
            const double gn = sqrt(g2), inv = 1.0 / gn;
#pragma unroll
            for (int d = 0; d < DIM; d++) nrm[d] *= inv;
            const double scale = iface ? gn / sqrt(r2) : 1.0;          // |F^-T n_ref| at this point
#pragma unroll
            for (int i = 0; i < ND; i++) {
              double sv, gv[DIM];
              h1_shape_grad<DIM, ORDER>(i, lam, gl, sv, gv);
              stage[lane * ND + i] = make_double4(sv, gv[0], gv[1], DIM == 3 ? gv[DIM - 1] : 0.0);
              double dn = 0.0;
#pragma unroll
              for (int d = 0; d < DIM; d++) dn += gv[d] * nrm[d];
              dnq[lane * ND + i] = dn;
            }
            facq[lane] = adet * (r.w * sfac * scale);
            if (hp) hq[lane] = DIM == 2 ? sqrt(adet) : cbrt(adet);     // mesh_size at this point
          }
          __syncwarp();
          for (int q = 0; q < nq; q++) {
            const double f = facq[q];
            double cm = cm0, cs = cs0, cn = cn0, cd = cd0;
            if (hp) {
              const double h = hq[q];
              cm *= ipow(h, A.form.p[0]); cs *= ipow(h, A.form.p[1]); cn *= ipow(h, A.form.p[2]); cd *= ipow(h, A.form.p[3]);
            }
#pragma unroll
            for (int t = 0; t < T; t++) {
              const double4 bi = stage[q * ND + ei[t]], bj = stage[q * ND + ej[t]];
              double gg = bi.y * bj.y + bi.z * bj.z + (DIM == 3 ? bi.w * bj.w : 0.0);
              double v = cs * gg + cm * bi.x * bj.x;
              if (nterm) {
                const double di = dnq[q * ND + ei[t]], dj = dnq[q * ND + ej[t]];
                v += cn * (dj * bi.x + di * bj.x) + cd * di * dj;
              }
              acc[t] += f * v;
            }
          }
          __syncwarp();
        }
      }
    }
#pragma unroll
    for (int t = 0; t < T; t++) {
      int idx = lane + 32 * t;
      if (idx < NE) A.elmats[k * NE + idx] = acc[t];
    }
    if (A.npts && lane == 0) A.npts[k] = np_used;
  }
}

// Integrate on the deformed mesh: warp per element, lane = rule point, fixed butterfly -> partial[e]
template <int DIM, int DORD>
__global__ void __launch_bounds__(DW * 32) k_int_def(DefArgs A, int64_t ne) {
  constexpr int NDG = H1<DIM, DORD>::ND;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *gc = reinterpret_cast<double *>(smem_raw) + warp * NDG * DIM;
  const int64_t nwarps = (int64_t)gridDim.x * DW;
  for (int64_t e = (int64_t)blockIdx.x * DW + warp; e < ne; e += nwarps) {
    const int32_t *ev = A.M.el2v + e * (DIM + 1);
    double ls[4];
#pragma unroll
    for (int k = 0; k <= DIM; k++) ls[k] = __ldg(A.lset + __ldg(ev + k));
    int mask;
    const int dom = sign_test<DIM>(ls, mask);
    const bool cut = dom == IF;
    double sum = 0.0;
    if (cut || dom == A.dt) {                                            // (warp-uniform)
      load_iso<DIM, DORD>(A, e, lane, ls, gc);
      double r2 = 0.0;
#pragma unroll
      for (int kk = 0; kk < DIM; kk++) { double v = ls[kk] - ls[DIM]; r2 += v * v; }
      const bool iface = cut && A.dt == IF;
      CutEntry E;
      if (cut) E = A.tab->e[DIM - 2][A.dt][mask]; else { E.ns = 1; E.npt = 0; }
      const double4 *rule = iface ? A.rule_if : A.rule_vol;
      const int nr = iface ? A.nr_if : A.nr_vol;
      for (int s = 0; s < E.ns; s++) {
        double P[4][3];
        double sfac = 1.0;
        if (cut) {
          for (int v = 0; v < E.npt; v++) coded_point<DIM>(E.v[s][v], ls, P[v]);
          sfac = simplex_factor(E.npt, P);
        }
        for (int q = lane; q < nr; q += 32) {
          double4 r = rule[q];
          double xi[3] = {r.x, r.y, r.z};
          if (cut) simplex_point(E.npt, P, r, xi);
          double lam[4] = {xi[0], xi[1], xi[2], 0.0};
          double last = 1.0;
#pragma unroll
          for (int kk = 0; kk < DIM; kk++) last -= lam[kk];
          lam[DIM] = last;
          double x[3] = {0.0, 0.0, 0.0}, gl[DIM + 1][DIM], det;
          iso_point<DIM, DORD>(GcWarp{gc}, lam, x, gl, det);
          double scale = 1.0;
          if (iface) {
            double g2 = 0.0;
#pragma unroll
            for (int d = 0; d < DIM; d++) {
              double v = 0.0;
#pragma unroll
              for (int kk = 0; kk <= DIM; kk++) v += ls[kk] * gl[kk][d];
              g2 += v * v;
            }
            scale = sqrt(g2) / sqrt(r2);
          }
          double fv = 0.0;
          for (int m = 0; m < A.f.n; m++)
            fv += A.f.c[m] * ipow(x[0], A.f.ex[m][0]) * ipow(x[1], A.f.ex[m][1]) * ipow(x[2], A.f.ex[m][2]);
          sum += fabs(det) * (r.w * sfac * scale) * fv;
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) A.partial[e] = sum;
  }
}

// ---------------------------------------------------------------------------------------------------------------
// THREAD PER ELEMENT (the production path).  The warp-per-element kernels above spend a warp instruction per rule point
// with 14 of 32 lanes busy and then 100 entries x 14 points on all lanes: ~1300 FP64 warp instructions per element.  One
// thread per element keeps the upper triangle (ND (ND+1)/2 accumulators) in registers and the geometry coefficients in
// its shared-memory column: ~7 K thread instructions = ~220 warp instructions per element.  Cut elements (14..42 points
// instead of 14) would stall nearly every warp (8 % of the list, 93 % of the warps), so the first pass only records them
// in a work list (integer atomics: the order never reaches a value) and a second pass of the same code walks that list.
// ---------------------------------------------------------------------------------------------------------------
constexpr int TT = 128;           // threads per CTA of the thread-per-element kernels

template <int DIM>
__device__ __forceinline__ void load_vertices(const ngsx_mesh &M, int64_t e, int *v) {
  if constexpr (DIM == 3) {
    int4 q = __ldg(reinterpret_cast<const int4 *>(M.el2v) + e);
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  } else {
    const int32_t *q = M.el2v + 3 * e;
    v[0] = __ldg(q); v[1] = __ldg(q + 1); v[2] = __ldg(q + 2);
  }
}

template <int DIM, int DORD>
__device__ __forceinline__ void load_gc_thread(const DefArgs &A, int64_t e, const int *v, double *gcs) {
  constexpr int NDG = H1<DIM, DORD>::ND;
#pragma unroll
  for (int i = 0; i < NDG; i++) {
    const int64_t g = __ldg(A.D.el2dof + e * NDG + i);
#pragma unroll
    for (int d = 0; d < DIM; d++) {
      double c = __ldg(A.D.coef + g * DIM + d);
      if (i <= DIM) c += __ldg(A.M.vert + (int64_t)v[i <= DIM ? i : 0] * DIM + d);
      gcs[(i * DIM + d) * TT] = c;
    }
  }
}

template <int DIM, int ORDER, int DORD, bool CUTPASS>
__global__ void __launch_bounds__(TT) k_elmat_def_t(DefArgs A, int64_t count) {
  constexpr int ND = H1<DIM, ORDER>::ND, NE = ND * ND, NDG = H1<DIM, DORD>::ND, NT = ND * (ND + 1) / 2;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *gcs = reinterpret_cast<double *>(smem_raw) + threadIdx.x;
  double4 *srule_vol = reinterpret_cast<double4 *>(smem_raw + (size_t)NDG * DIM * TT * 8);
  double4 *srule_if = srule_vol + A.nr_vol;
  for (int i = threadIdx.x; i < A.nr_vol; i += TT) srule_vol[i] = A.rule_vol[i];
  for (int i = threadIdx.x; i < A.nr_if; i += TT) srule_if[i] = A.rule_if[i];
  __syncthreads();
  const bool hp = (A.form.p[0] | A.form.p[1] | A.form.p[2] | A.form.p[3]) != 0;
  const bool nterm = A.form.c[2] != 0.0 || A.form.c[3] != 0.0;
  const int64_t n = CUTPASS ? (int64_t)*A.wcount : count;
  for (int64_t w = (int64_t)blockIdx.x * TT + threadIdx.x; w < n; w += (int64_t)gridDim.x * TT) {
    const int64_t k = CUTPASS ? (int64_t)A.worklist[w] : w;
    const int64_t e = A.els ? A.els[k] : k;
    int v[4];
    load_vertices<DIM>(A.M, e, v);
    double ls[4];
#pragma unroll
    for (int a = 0; a <= DIM; a++) ls[a] = __ldg(A.lset + v[a]);
    int mask;
    const int dom = sign_test<DIM>(ls, mask);
    const bool cut = dom == IF;
    double *out = A.elmats + k * NE;
    if constexpr (!CUTPASS) {
      if (dom < 0 && A.err) {
        atomicAdd(&A.err[0], 1ULL);
        atomicMin(reinterpret_cast<long long *>(&A.err[1]), (long long)e);
      }
      if (cut) {                                                        // second pass
        A.worklist[atomicAdd(A.wcount, 1)] = (int32_t)k;
        continue;
      }
      if (dom != A.dt) {
        if (A.rowslot) {
          constexpr int NDP = ND + (ND & 1);
          for (int i = 0; i < ND; i++) {
            const int slot = __ldg(A.rowslot + k * ND + i);
            if (slot < 0) continue;
            double *o = A.elmats + (int64_t)slot * NDP;
#pragma unroll
            for (int j = 0; j < NDP; j++) o[j] = 0.0;
          }
        } else {
#pragma unroll 4
          for (int t = 0; t < NE; t++) out[t] = 0.0;
        }
        if (A.npts) A.npts[k] = 0;
        continue;
      }
    }
    load_gc_thread<DIM, DORD>(A, e, v, gcs);
    double acc[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) acc[t] = 0.0;
    double r2 = 0.0;
#pragma unroll
    for (int a = 0; a < DIM; a++) { double d = ls[a] - ls[DIM]; r2 += d * d; }
    const bool iface = CUTPASS && A.dt == IF;
    CutEntry E;
    if (CUTPASS) E = A.tab->e[DIM - 2][A.dt][mask]; else { E.ns = 1; E.npt = 0; }
    const double4 *rule = iface ? srule_if : srule_vol;
    const int nr = iface ? A.nr_if : A.nr_vol;
    for (int s = 0; s < E.ns; s++) {
      double P[4][3];
      double sfac = 1.0;
      if (CUTPASS) {
        for (int a = 0; a < E.npt; a++) coded_point<DIM>(E.v[s][a], ls, P[a]);
        sfac = simplex_factor(E.npt, P);
      }
      for (int q = 0; q < nr; q++) {
        const double4 r = rule[q];
        double xi[3] = {r.x, r.y, r.z};
        if (CUTPASS) simplex_point(E.npt, P, r, xi);
        double lam[4] = {xi[0], xi[1], xi[2], 0.0};
        double last = 1.0;
#pragma unroll
        for (int a = 0; a < DIM; a++) last -= lam[a];
        lam[DIM] = last;
        double x[DIM], gl[DIM + 1][DIM], det;
        iso_point<DIM, DORD>(GcThread<TT>{gcs}, lam, x, gl, det);
        const double adet = fabs(det);
        double cm = A.form.c[0], cs = A.form.c[1], cn = A.form.c[2], cd = A.form.c[3];
        if (hp) {                                                       // mesh_size at this point (not elementwise constant)
          const double h = DIM == 2 ? sqrt(adet) : cbrt(adet);
          cm *= ipow(h, A.form.p[0]); cs *= ipow(h, A.form.p[1]); cn *= ipow(h, A.form.p[2]); cd *= ipow(h, A.form.p[3]);
        }
        double nrm[DIM], g2 = 0.0;
#pragma unroll
        for (int d = 0; d < DIM; d++) {
          double t = 0.0;
#pragma unroll
          for (int a = 0; a <= DIM; a++) t += ls[a] * gl[a][d];
          nrm[d] = t; g2 += t * t;
        }
        const double gn = sqrt(g2);
        const double f = adet * (r.w * sfac * (iface ? gn / sqrt(r2) : 1.0));      // |det F| w |F^-T n_ref|
        double S[ND], G[ND][DIM];
#pragma unroll
        for (int i = 0; i < ND; i++) h1_shape_grad<DIM, ORDER>(i, lam, gl, S[i], G[i]);
        const double fs = f * cs, fm = f * cm;
        if (!nterm) {
          int t = 0;
#pragma unroll
          for (int i = 0; i < ND; i++) {
            double gi[DIM];
#pragma unroll
            for (int d = 0; d < DIM; d++) gi[d] = fs * G[i][d];
            const double si = fm * S[i];
#pragma unroll
            for (int j = 0; j <= i; j++) {
              double a = acc[t];
#pragma unroll
              for (int d = 0; d < DIM; d++) a = fma(gi[d], G[j][d], a);
              acc[t++] = fma(si, S[j], a);
            }
          }
        } else {
          const double inv = 1.0 / gn, fn = f * cn, fd = f * cd;
          double DN[ND];
#pragma unroll
          for (int i = 0; i < ND; i++) {
            double t = 0.0;
#pragma unroll
            for (int d = 0; d < DIM; d++) t += G[i][d] * nrm[d];
            DN[i] = t * inv;
          }
          int t = 0;
#pragma unroll
          for (int i = 0; i < ND; i++) {
            double gi[DIM];
#pragma unroll
            for (int d = 0; d < DIM; d++) gi[d] = fs * G[i][d];
            const double si = fm * S[i] + fn * DN[i];                   // multiplies S_j
            const double di = fn * S[i] + fd * DN[i];                   // multiplies DN_j
#pragma unroll
            for (int j = 0; j <= i; j++) {
              double a = acc[t];
#pragma unroll
              for (int d = 0; d < DIM; d++) a = fma(gi[d], G[j][d], a);
              a = fma(si, S[j], a);
              acc[t++] = fma(di, DN[j], a);
            }
          }
        }
      }
    }
    {                                                                   // mirrored store of the symmetric block
      auto tri = [](int i, int j) { return i >= j ? i * (i + 1) / 2 + j : j * (j + 1) / 2 + i; };
      if (A.rowslot) {                                                  // packed rows: row i to the slot of the pair (k, i)
        constexpr int NDP = ND + (ND & 1);
#pragma unroll
        for (int i = 0; i < ND; i++) {
          const int slot = __ldg(A.rowslot + k * ND + i);
          if (slot >= 0) {
            double2 *o2 = reinterpret_cast<double2 *>(A.elmats + (int64_t)slot * NDP);
#pragma unroll
            for (int j = 0; j < NDP / 2; j++)
              o2[j] = make_double2(acc[tri(i, 2 * j)], 2 * j + 1 < ND ? acc[tri(i, 2 * j + 1)] : 0.0);
          }
        }
      } else if constexpr (NE % 2 == 0) {                                      // 16-byte stores: the block starts 16-byte aligned (NE even)
        double2 *o2 = reinterpret_cast<double2 *>(out);
#pragma unroll
        for (int t = 0; t < NE / 2; t++)
          o2[t] = make_double2(acc[tri((2 * t) / ND, (2 * t) % ND)], acc[tri((2 * t + 1) / ND, (2 * t + 1) % ND)]);
      } else {
#pragma unroll
        for (int t = 0; t < NE; t++) out[t] = acc[tri(t / ND, t % ND)];
      }
    }
    if (A.npts) A.npts[k] = E.ns * nr;
  }
}

// element vectors (SymbolicCutLFI on a deformed mesh, xfem/symboliccutlfi.cpp:52-158): thread per element, cut elements inline
// (per point only the ND shape values and the integrand at the mapped point: the quadrature is a small part of the work)
template <int DIM, int ORDER, int DORD>
__global__ void __launch_bounds__(TT) k_elvec_def_t(DefArgs A, int64_t count) {
  constexpr int ND = H1<DIM, ORDER>::ND, NDG = H1<DIM, DORD>::ND;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *gcs = reinterpret_cast<double *>(smem_raw) + threadIdx.x;
  double4 *srule_vol = reinterpret_cast<double4 *>(smem_raw + (size_t)NDG * DIM * TT * 8);
  double4 *srule_if = srule_vol + A.nr_vol;
  for (int i = threadIdx.x; i < A.nr_vol; i += TT) srule_vol[i] = A.rule_vol[i];
  for (int i = threadIdx.x; i < A.nr_if; i += TT) srule_if[i] = A.rule_if[i];
  __syncthreads();
  for (int64_t k = (int64_t)blockIdx.x * TT + threadIdx.x; k < count; k += (int64_t)gridDim.x * TT) {
    const int64_t e = A.els ? A.els[k] : k;
    int v[4];
    load_vertices<DIM>(A.M, e, v);
    double ls[4];
#pragma unroll
    for (int a = 0; a <= DIM; a++) ls[a] = __ldg(A.lset + v[a]);
    int mask;
    const int dom = sign_test<DIM>(ls, mask);
    const bool cut = dom == IF;
    if (dom < 0 && A.err) {
      atomicAdd(&A.err[0], 1ULL);
      atomicMin(reinterpret_cast<long long *>(&A.err[1]), (long long)e);
    }
    double acc[ND];
#pragma unroll
    for (int i = 0; i < ND; i++) acc[i] = 0.0;
    int np_used = 0;
    if (cut || dom == A.dt) {
      load_gc_thread<DIM, DORD>(A, e, v, gcs);
      double r2 = 0.0;
#pragma unroll
      for (int a = 0; a < DIM; a++) { double d = ls[a] - ls[DIM]; r2 += d * d; }
      const bool iface = cut && A.dt == IF;
      CutEntry E;
      if (cut) E = A.tab->e[DIM - 2][A.dt][mask]; else { E.ns = 1; E.npt = 0; }
      const double4 *rule = iface ? srule_if : srule_vol;
      const int nr = iface ? A.nr_if : A.nr_vol;
      np_used = E.ns * nr;
      for (int s = 0; s < E.ns; s++) {
        double P[4][3];
        double sfac = 1.0;
        if (cut) {
          for (int a = 0; a < E.npt; a++) coded_point<DIM>(E.v[s][a], ls, P[a]);
          sfac = simplex_factor(E.npt, P);
        }
        for (int q = 0; q < nr; q++) {
          const double4 r = rule[q];
          double xi[3] = {r.x, r.y, r.z};
          if (cut) simplex_point(E.npt, P, r, xi);
          double lam[4] = {xi[0], xi[1], xi[2], 0.0};
          double last = 1.0;
#pragma unroll
          for (int a = 0; a < DIM; a++) last -= lam[a];
          lam[DIM] = last;
          double x[3] = {0.0, 0.0, 0.0}, gl[DIM + 1][DIM], det;
          iso_point<DIM, DORD>(GcThread<TT>{gcs}, lam, x, gl, det);
          double scale = 1.0;
          if (iface) {
            double g2 = 0.0;
#pragma unroll
            for (int d = 0; d < DIM; d++) {
              double t = 0.0;
#pragma unroll
              for (int a = 0; a <= DIM; a++) t += ls[a] * gl[a][d];
              g2 += t * t;
            }
            scale = sqrt(g2) / sqrt(r2);
          }
          double fv = 0.0;
          for (int m = 0; m < A.f.n; m++)
            fv += A.f.c[m] * ipow(x[0], A.f.ex[m][0]) * ipow(x[1], A.f.ex[m][1]) * ipow(x[2], A.f.ex[m][2]);
          const double pv = fabs(det) * (r.w * sfac * scale) * fv;
#pragma unroll
          for (int i = 0; i < ND; i++) acc[i] += pv * h1_shape<DIM, ORDER>(i, lam);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < ND; i++) A.elmats[k * ND + i] = acc[i];
    if (A.npts) A.npts[k] = np_used;
  }
}

// Integrate, thread per element, two passes like the element matrices: pass 1 integrates the uncut elements of the domain into
// val[e] and records the cut ones; pass 2 walks the recorded list and writes val[e] of the cut elements.  (Inline, a cut element --
// 14..42 points -- held up its whole warp: 149 K cut elements touch more warps than the 59 K that carry NEG elements; measured
// 1.28 ms inline.)  The values sit at their element's position, so the order of the work list never reaches the sum: the
// fixed-shape tree over val[0..ne) is bitwise reproducible.
template <int DIM, int DORD, bool CUTPASS>
__global__ void __launch_bounds__(TT) k_int_def_t(DefArgs A, int64_t ne) {
  constexpr int NDG = H1<DIM, DORD>::ND;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *gcs = reinterpret_cast<double *>(smem_raw) + threadIdx.x;
  double4 *srule_vol = reinterpret_cast<double4 *>(smem_raw + (size_t)NDG * DIM * TT * 8);
  double4 *srule_if = srule_vol + A.nr_vol;
  for (int i = threadIdx.x; i < A.nr_vol; i += TT) srule_vol[i] = A.rule_vol[i];
  for (int i = threadIdx.x; i < A.nr_if; i += TT) srule_if[i] = A.rule_if[i];
  __syncthreads();
  const int64_t n = CUTPASS ? (int64_t)*A.wcount : ne;
  for (int64_t w = (int64_t)blockIdx.x * TT + threadIdx.x; w < n; w += (int64_t)gridDim.x * TT) {
    const int64_t e = CUTPASS ? (int64_t)A.worklist[w] : w;
    int v[4];
    load_vertices<DIM>(A.M, e, v);
    double ls[4];
#pragma unroll
    for (int a = 0; a <= DIM; a++) ls[a] = __ldg(A.lset + v[a]);
    int mask;
    const int dom = sign_test<DIM>(ls, mask);
    const bool cut = dom == IF;
    if constexpr (!CUTPASS) {
      if (cut) A.worklist[atomicAdd(A.wcount, 1)] = (int32_t)e;
      if (cut || dom != A.dt) { A.partial[e] = 0.0; continue; }
    }
    double sum = 0.0;
    load_gc_thread<DIM, DORD>(A, e, v, gcs);
    double r2 = 0.0;
#pragma unroll
    for (int a = 0; a < DIM; a++) { double d = ls[a] - ls[DIM]; r2 += d * d; }
    const bool iface = CUTPASS && A.dt == IF;
    CutEntry E;
    if (CUTPASS) E = A.tab->e[DIM - 2][A.dt][mask]; else { E.ns = 1; E.npt = 0; }
    const double4 *rule = iface ? srule_if : srule_vol;
    const int nr = iface ? A.nr_if : A.nr_vol;
    for (int s = 0; s < E.ns; s++) {
      double P[4][3];
      double sfac = 1.0;
      if (CUTPASS) {
        for (int a = 0; a < E.npt; a++) coded_point<DIM>(E.v[s][a], ls, P[a]);
        sfac = simplex_factor(E.npt, P);
      }
      for (int q = 0; q < nr; q++) {
        const double4 r = rule[q];
        double xi[3] = {r.x, r.y, r.z};
        if (CUTPASS) simplex_point(E.npt, P, r, xi);
        double lam[4] = {xi[0], xi[1], xi[2], 0.0};
        double last = 1.0;
#pragma unroll
        for (int a = 0; a < DIM; a++) last -= lam[a];
        lam[DIM] = last;
        double x[3] = {0.0, 0.0, 0.0}, gl[DIM + 1][DIM], det;
        iso_point<DIM, DORD>(GcThread<TT>{gcs}, lam, x, gl, det);
        double scale = 1.0;
        if (iface) {
          double g2 = 0.0;
#pragma unroll
          for (int d = 0; d < DIM; d++) {
            double t = 0.0;
#pragma unroll
            for (int a = 0; a <= DIM; a++) t += ls[a] * gl[a][d];
            g2 += t * t;
          }
          scale = sqrt(g2) / sqrt(r2);
        }
        double fv = 0.0;
        for (int m = 0; m < A.f.n; m++)
          fv += A.f.c[m] * ipow(x[0], A.f.ex[m][0]) * ipow(x[1], A.f.ex[m][1]) * ipow(x[2], A.f.ex[m][2]);
        sum += fabs(det) * (r.w * sfac * scale) * fv;
      }
    }
    A.partial[e] = sum;
  }
}

// fixed-shape reduction (device independent: the shape depends on n only)
__global__ void __launch_bounds__(256) k_def_reduce(int64_t n, const double *__restrict__ in, double *__restrict__ partial) {
  __shared__ double wsum[8];
  double sum = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) sum += in[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < 8; w++) s += wsum[w];
    partial[blockIdx.x] = s;
  }
}
__global__ void __launch_bounds__(1024) k_def_reduce_final(int64_t n, const double *__restrict__ partial, double *__restrict__ out) {
  __shared__ double s[1024];
  double acc = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += 1024) acc += partial[i];
  s[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = s[0];
}

template <typename K>
static int grid_for(K kernel, size_t smem, int sm_count, int64_t nwarp_items, unsigned *grid) {
  NGSX_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, DW * 32, smem));
  if (occ < 1) { set_error("kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
  int64_t need = (nwarp_items + DW - 1) / DW;
  *grid = (unsigned)std::max<int64_t>(1, std::min(need, (int64_t)sm_count * occ));
  return NGSX_OK;
}

template <typename K>
static int grid_threads(K kernel, size_t smem, int sm_count, int64_t nitems, unsigned *grid) {
  NGSX_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, TT, smem));
  if (occ < 1) { set_error("kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
  *grid = (unsigned)std::max<int64_t>(1, std::min((nitems + TT - 1) / TT, (int64_t)sm_count * occ));
  return NGSX_OK;
}

static int g_def_warp = 0;      // debug switch: the first (warp-per-element) kernels

template <int DIM, int ORDER, int DORD>
static int launch_elmat_def(DefArgs &A, int sm_count, int64_t nlist, cudaStream_t st) {
  constexpr int ND = H1<DIM, ORDER>::ND, NDG = H1<DIM, DORD>::ND;
  if (g_def_warp) {
    const size_t smem = DW * ((size_t)DQ * ND * 32 + (size_t)DQ * ND * 8 + 2 * DQ * 8 + NDG * DIM * 8);
    auto kern = k_elmat_def<DIM, ORDER, DORD>;
    unsigned grid;
    if (int rc = grid_for(kern, smem, sm_count, nlist, &grid)) return rc;
    kern<<<grid, DW * 32, smem, st>>>(A, nlist);
    NGSX_LAUNCH_CHECK();
    return NGSX_OK;
  }
  // pass 1: uncut elements (cut ones recorded), pass 2: the recorded cut elements
  int32_t *work = nullptr;
  NGSX_CUDA(cudaMallocAsync(&work, (nlist + 1) * sizeof(int32_t), st));
  A.worklist = work + 1; A.wcount = work;
  cudaError_t e0 = cudaMemsetAsync(work, 0, sizeof(int32_t), st);
  const size_t smem = (size_t)NDG * DIM * TT * 8 + (size_t)(A.nr_vol + A.nr_if) * 32;
  auto k1 = k_elmat_def_t<DIM, ORDER, DORD, false>;
  auto k2 = k_elmat_def_t<DIM, ORDER, DORD, true>;
  unsigned g1 = 1, g2 = 1;
  int rc = grid_threads(k1, smem, sm_count, nlist, &g1);
  if (!rc) rc = grid_threads(k2, smem, sm_count, nlist, &g2);
  if (!rc) {
    k1<<<g1, TT, smem, st>>>(A, nlist);
    k2<<<g2, TT, smem, st>>>(A, nlist);
    g_launches.fetch_add(2);
  }
  cudaError_t e1 = cudaGetLastError();
  cudaFreeAsync(work, st);
  if (rc) return rc;
  NGSX_CUDA(e0);
  NGSX_CUDA(e1);
  return NGSX_OK;
}

template <int DIM, int ORDER, int DORD>
static int launch_elvec_def(const DefArgs &A, int sm_count, int64_t nlist, cudaStream_t st) {
  constexpr int NDG = H1<DIM, DORD>::ND;
  const size_t smem = (size_t)NDG * DIM * TT * 8 + (size_t)(A.nr_vol + A.nr_if) * 32;
  auto kern = k_elvec_def_t<DIM, ORDER, DORD>;
  unsigned grid;
  if (int rc = grid_threads(kern, smem, sm_count, nlist, &grid)) return rc;
  kern<<<grid, TT, smem, st>>>(A, nlist);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

template <int DIM, int DORD>
static int launch_int_def(DefArgs &A, int sm_count, int64_t ne, cudaStream_t st) {
  constexpr int NDG = H1<DIM, DORD>::ND;
  if (g_def_warp) {
    const size_t smem = DW * (size_t)NDG * DIM * 8;
    auto kern = k_int_def<DIM, DORD>;
    unsigned grid;
    if (int rc = grid_for(kern, smem, sm_count, ne, &grid)) return rc;
    kern<<<grid, DW * 32, smem, st>>>(A, ne);
    NGSX_LAUNCH_CHECK();
    return NGSX_OK;
  }
  int32_t *work = nullptr;
  NGSX_CUDA(cudaMallocAsync(&work, (ne + 1) * sizeof(int32_t), st));
  A.worklist = work + 1; A.wcount = work;
  cudaError_t e0 = cudaMemsetAsync(work, 0, sizeof(int32_t), st);
  const size_t smem = (size_t)NDG * DIM * TT * 8 + (size_t)(A.nr_vol + A.nr_if) * 32;
  auto k1 = k_int_def_t<DIM, DORD, false>;
  auto k2 = k_int_def_t<DIM, DORD, true>;
  unsigned g1 = 1, g2 = 1;
  int rc = grid_threads(k1, smem, sm_count, ne, &g1);
  if (!rc) rc = grid_threads(k2, smem, sm_count, ne, &g2);
  if (!rc) {
    k1<<<g1, TT, smem, st>>>(A, ne);
    k2<<<g2, TT, smem, st>>>(A, ne);
    g_launches.fetch_add(2);
  }
  cudaError_t e1 = cudaGetLastError();
  cudaFreeAsync(work, st);
  if (rc) return rc;
  NGSX_CUDA(e0);
  NGSX_CUDA(e1);
  return NGSX_OK;
}

static int check_common(const ngsx_mesh *mesh, const ngsx_deformation *defo, const double *lset_v, int dt, int order,
                        const DeviceData **dd, DefArgs &A) {
  NGSX_REQUIRE(mesh && (mesh->dim == 2 || mesh->dim == 3), "mesh.dim must be 2 or 3");
  NGSX_REQUIRE(lset_v && mesh->vert && mesh->el2v, "mesh / level-set arrays are NULL");
  NGSX_REQUIRE(dt >= 0 && dt <= 2, "domain_type must be NEG, POS or IF");
  NGSX_REQUIRE(defo && defo->el2dof && defo->coef, "deformation arrays are NULL");
  if (defo->order < 1 || defo->order > 2) {
    set_error("deformation order outside the CUDA path (supported: 1, 2); no CPU fallback");
    return NGSX_ERR_UNSUPPORTED;
  }
  const int ndg = defo->order == 1 ? mesh->dim + 1 : (mesh->dim == 2 ? 6 : 10);
  NGSX_REQUIRE(defo->nd == ndg, "deformation.nd does not match the H1 element of that order");
  *dd = device_data();
  if (!*dd) return NGSX_ERR_CUDA;
  if (!rule_lookup(*dd, mesh->dim, order, &A.rule_vol, &A.nr_vol) || !rule_lookup(*dd, mesh->dim - 1, order, &A.rule_if, &A.nr_if)) {
    set_error("integration order beyond the tabulated reference rules");
    return NGSX_ERR_UNSUPPORTED;
  }
  A.M = *mesh; A.D.el2dof = defo->el2dof; A.D.coef = defo->coef; A.lset = lset_v; A.tab = (*dd)->cut_table; A.dt = dt;
  return NGSX_OK;
}

}  // namespace ngsx

using namespace ngsx;

extern "C" void ngsx_debug_set_deformed_warp(int on) { g_def_warp = on; }

extern "C" int ngsx_elmat_h1_deformed(const ngsx_mesh *mesh, const ngsx_deformation *defo, const double *lset_v, int fe_order, int dt,
                                      int intorder, const ngsx_form_h1 *h_form, int64_t nlist, const int32_t *els, double *elmats,
                                      int32_t *npts, int64_t *err, void *stream) {
  NGSX_REQUIRE(h_form && (nlist == 0 || elmats) && nlist >= 0, "elmat_h1_deformed: bad arguments");
  if (fe_order < 1 || fe_order > 2) {
    set_error("H1 order outside the CUDA path (supported: 1, 2); no CPU fallback");
    return NGSX_ERR_UNSUPPORTED;
  }
  const DeviceData *dd;
  DefArgs A{};
  if (int rc = check_common(mesh, defo, lset_v, dt, intorder, &dd, A)) return rc;
  cudaStream_t st = as_stream(stream);
  if (err) { if (int rc = reset_err(err, st)) return rc; }
  if (nlist == 0) return NGSX_OK;
  A.form.c[0] = h_form->c_mass; A.form.c[1] = h_form->c_stiff; A.form.c[2] = h_form->c_nitsche; A.form.c[3] = h_form->c_dndn;
  A.form.p[0] = h_form->hpow_mass; A.form.p[1] = h_form->hpow_stiff; A.form.p[2] = h_form->hpow_nitsche; A.form.p[3] = h_form->hpow_dndn;
  A.els = els; A.elmats = elmats; A.npts = npts; A.err = reinterpret_cast<unsigned long long *>(err);
  A.rowslot = take_packed_output();
  if (A.rowslot && g_def_warp) { set_error("elmat_h1_deformed: packed output is served by the thread-per-element kernels only"); return NGSX_ERR_UNSUPPORTED; }
  switch (mesh->dim * 100 + fe_order * 10 + defo->order) {
    case 211: return launch_elmat_def<2, 1, 1>(A, dd->sm_count, nlist, st);
    case 212: return launch_elmat_def<2, 1, 2>(A, dd->sm_count, nlist, st);
    case 221: return launch_elmat_def<2, 2, 1>(A, dd->sm_count, nlist, st);
    case 222: return launch_elmat_def<2, 2, 2>(A, dd->sm_count, nlist, st);
    case 311: return launch_elmat_def<3, 1, 1>(A, dd->sm_count, nlist, st);
    case 312: return launch_elmat_def<3, 1, 2>(A, dd->sm_count, nlist, st);
    case 321: return launch_elmat_def<3, 2, 1>(A, dd->sm_count, nlist, st);
    case 322: return launch_elmat_def<3, 2, 2>(A, dd->sm_count, nlist, st);
  }
  set_error("unsupported (dim, order, deformation order)");
  return NGSX_ERR_UNSUPPORTED;
}

extern "C" int ngsx_elvec_h1_deformed(const ngsx_mesh *mesh, const ngsx_deformation *defo, const double *lset_v, int fe_order, int dt,
                                      int intorder, const ngsx_poly *h_f, int64_t nlist, const int32_t *els, double *elvecs,
                                      int32_t *npts, int64_t *err, void *stream) {
  NGSX_REQUIRE(h_f && (nlist == 0 || elvecs) && nlist >= 0, "elvec_h1_deformed: bad arguments");
  NGSX_REQUIRE(h_f->nterms >= 0 && h_f->nterms <= NGSX_POLY_MAXTERMS, "elvec_h1_deformed: too many polynomial terms");
  if (fe_order < 1 || fe_order > 2) {
    set_error("H1 order outside the CUDA path (supported: 1, 2); no CPU fallback");
    return NGSX_ERR_UNSUPPORTED;
  }
  const DeviceData *dd;
  DefArgs A{};
  if (int rc = check_common(mesh, defo, lset_v, dt, intorder, &dd, A)) return rc;
  cudaStream_t st = as_stream(stream);
  if (err) { if (int rc = reset_err(err, st)) return rc; }
  if (nlist == 0) return NGSX_OK;
  A.f.n = h_f->nterms;
  for (int m = 0; m < A.f.n; m++) { A.f.c[m] = h_f->coef[m]; for (int d = 0; d < 3; d++) A.f.ex[m][d] = h_f->ex[m][d]; }
  A.els = els; A.elmats = elvecs; A.npts = npts; A.err = reinterpret_cast<unsigned long long *>(err);
  switch (mesh->dim * 100 + fe_order * 10 + defo->order) {
    case 211: return launch_elvec_def<2, 1, 1>(A, dd->sm_count, nlist, st);
    case 212: return launch_elvec_def<2, 1, 2>(A, dd->sm_count, nlist, st);
    case 221: return launch_elvec_def<2, 2, 1>(A, dd->sm_count, nlist, st);
    case 222: return launch_elvec_def<2, 2, 2>(A, dd->sm_count, nlist, st);
    case 311: return launch_elvec_def<3, 1, 1>(A, dd->sm_count, nlist, st);
    case 312: return launch_elvec_def<3, 1, 2>(A, dd->sm_count, nlist, st);
    case 321: return launch_elvec_def<3, 2, 1>(A, dd->sm_count, nlist, st);
    case 322: return launch_elvec_def<3, 2, 2>(A, dd->sm_count, nlist, st);
  }
  set_error("unsupported (dim, order, deformation order)");
  return NGSX_ERR_UNSUPPORTED;
}

extern "C" int ngsx_integrate_deformed(const ngsx_mesh *mesh, const ngsx_deformation *defo, const double *lset_v, int dt, int order,
                                       const ngsx_poly *h_f, double *out, void *stream) {
  NGSX_REQUIRE(h_f && out, "integrate_deformed: bad arguments");
  NGSX_REQUIRE(h_f->nterms >= 0 && h_f->nterms <= NGSX_POLY_MAXTERMS, "integrate_deformed: too many polynomial terms");
  const DeviceData *dd;
  DefArgs A{};
  if (int rc = check_common(mesh, defo, lset_v, dt, order, &dd, A)) return rc;
  cudaStream_t st = as_stream(stream);
  const int64_t ne = mesh->ne;
  if (ne == 0) { NGSX_CUDA(cudaMemsetAsync(out, 0, sizeof(double), st)); return NGSX_OK; }
  A.f.n = h_f->nterms;
  for (int m = 0; m < A.f.n; m++) { A.f.c[m] = h_f->coef[m]; for (int d = 0; d < 3; d++) A.f.ex[m][d] = h_f->ex[m][d]; }
  const int64_t npart = ne;                                                      // one value per element, at the element's position
  const unsigned nmid = (unsigned)std::min<int64_t>((npart + 255) / 256, 296);   // fixed upper bound: device-independent shape
  double *partial = nullptr;
  NGSX_CUDA(cudaMallocAsync(&partial, (npart + nmid) * sizeof(double), st));
  A.partial = partial;
  int rc;
  switch (mesh->dim * 10 + defo->order) {
    case 21: rc = launch_int_def<2, 1>(A, dd->sm_count, ne, st); break;
    case 22: rc = launch_int_def<2, 2>(A, dd->sm_count, ne, st); break;
    case 31: rc = launch_int_def<3, 1>(A, dd->sm_count, ne, st); break;
    default: rc = launch_int_def<3, 2>(A, dd->sm_count, ne, st); break;
  }
  if (rc == NGSX_OK) {
    k_def_reduce<<<nmid, 256, 0, st>>>(npart, partial, partial + npart);
    k_def_reduce_final<<<1, 1024, 0, st>>>(nmid, partial + npart, out);
    g_launches.fetch_add(2);
  }
  cudaError_t e1 = cudaGetLastError();
  cudaFreeAsync(partial, st);
  if (rc) return rc;
  NGSX_CUDA(e1);
  return NGSX_OK;
}
