// ngsx_xfem.cu -- (f-2) cut element matrices on the compound space [Vh, XFESpace(Vh)] with the sign-masked operators
// neg / pos / neg_grad / pos_grad of the reference (xfem/xfemdiffops.cpp:8-66: row i of the B-matrix of an x-operator is the
// base shape (gradient) where the sign of x-dof i equals the operator's domain, zero elsewhere; signs = XFESpace::domofdof,
// xfem/xFESpace.cpp:346-366) inside SymbolicCutBilinearFormIntegrator::T_CalcElementMatrixAdd (xfem/symboliccutbfi.cpp:73-315).
//
// The bilinear form arrives as a list of product terms  c * h^p * kappa * (L_v v)(L_u u)  (ngsx_xterm).  Every such term is
// a masked copy of one of five BASE matrices of the plain H1 shapes over the cut rule of the element,
//   M_ij = int phi_i phi_j,  N_ij = int phi_i dn phi_j,  D_ij = int dn phi_i dn phi_j,  K_ij = int grad phi_i . grad phi_j,
// (n = grad lset / |grad lset|, elementwise constant for a P1 level set), so the kernel integrates the base matrices once and
// then forms the (2 nd) x (2 nd) compound block  [std dofs | x dofs]  by mask algebra.
// The blocks are scattered by the owner-computes gather (ngsx_assemble_csr_rows_mats); several integrators of one form sum
// their blocks in one buffer (accumulate) and share one gather.
#include <algorithm>
#include "ngsx_host.cuh"
#include "ngsx_iso.cuh"

namespace ngsx {

constexpr int XMAXTERMS = 32;

struct XTermsDev {
  int n;
  double coef[XMAXTERMS];
  int8_t hpow[XMAXTERMS], kappa[XMAXTERMS], pair[XMAXTERMS], blk_u[XMAXTERMS], blk_v[XMAXTERMS];
  int8_t cls[10];             // terms are sorted by class blk_v * 3 + blk_u; class c = terms [cls[c], cls[c + 1])
};

template <int DIM>
__device__ __forceinline__ void xload_el(const ngsx_mesh &M, const double *__restrict__ lset, int64_t e, double *ls, double (*p)[DIM]) {
#pragma unroll
  for (int k = 0; k <= DIM; k++) {
    const int v = __ldg(M.el2v + e * (DIM + 1) + k);
    ls[k] = __ldg(lset + v);
#pragma unroll
    for (int d = 0; d < DIM; d++) p[k][d] = __ldg(M.vert + (int64_t)v * DIM + d);
  }
}

__device__ __forceinline__ double xipow(double x, int p) {
  double r = 1.0;
  if (p < 0) { x = 1.0 / x; p = -p; }
  for (int k = 0; k < p; k++) r *= x;
  return r;
}

// ---------------------------------------------------------------------------------------------------------------
// ONE THREAD PER ELEMENT, from monomial moments (as k_cutmat_mom, ngsx_assemble.cu): the hierarchical
// shapes are lam_i and -1/2 lam_a lam_b, so all four base matrices are fixed combinations of the moments int lam^alpha over
// the cut part (degree <= 4), of g_c = grad lam_c . n and of the Gram matrix grad lam_a . grad lam_b.  Per rule point the
// thread accumulates the 60 moments; per element it folds the form's terms into 5 coefficients per block class
// (shared memory, 45 doubles per thread), parks the moments in its shared-memory slice (runtime-indexed by the expansion
// loops: fully unrolled, the 400 entries of a 3D P2 block spilled 3.8 KB per thread) and writes the (2 nd)^2 block entry by
// entry.  The first version -- a warp per element staging shape values of 16 points through shared memory, lanes = entries --
// took 0.69-0.77 ms per integrator for 93 120 P2 tets; this one 0.38 ms (0.53 with accumulate), bound by its 4 warps per SM.
// ---------------------------------------------------------------------------------------------------------------
#ifndef NGSX_XMT
#define NGSX_XMT 128
#endif
constexpr int XMT = NGSX_XMT;     // threads per CTA of the moment kernel (120 doubles of shared memory per thread: one CTA per SM; 64 x 3 measured no better)
constexpr int XCO = 45;           // 9 classes x 5 operator pairs (odd stride: conflict-free)

// moments + geometry of one element in the thread's shared-memory slice (runtime-indexed by the expansion loops)
template <int DIM, int ORDER>
struct XMom {
  static constexpr int ND = H1<DIM, ORDER>::ND, NV = DIM + 1, NEDGE = ND - NV, N2 = NV * (NV + 1) / 2;
  static constexpr int NT = NEDGE > 0 ? NEDGE : 1, NQ = NEDGE > 0 ? NEDGE * (NEDGE + 1) / 2 : 1;
  static constexpr int O_M1 = 1, O_M2 = O_M1 + NV, O_MT = O_M2 + N2, O_MQ = O_MT + NV * NT, O_G = O_MQ + NQ, O_GM = O_G + NV;
  static constexpr int TOT = (O_GM + N2) | 1;                          // odd stride: conflict-free
  const double *w;
  __device__ __forceinline__ static int sym(int a, int b) { return a >= b ? a * (a + 1) / 2 + b : b * (b + 1) / 2 + a; }
  __device__ __forceinline__ double m0() const { return w[0]; }
  __device__ __forceinline__ double m1(int a) const { return w[O_M1 + a]; }
  __device__ __forceinline__ double m2(int a, int b) const { return w[O_M2 + sym(a, b)]; }
  __device__ __forceinline__ double mT(int v, int ed) const { return w[O_MT + v * NT + ed]; }
  __device__ __forceinline__ double mQ(int a, int b) const { return w[O_MQ + sym(a, b)]; }
  __device__ __forceinline__ double g(int a) const { return w[O_G + a]; }
  __device__ __forceinline__ double gg(int a, int b) const { return w[O_GM + sym(a, b)]; }
  // int phi_i phi_j
  __device__ __forceinline__ double M(int i, int j) const {
    if (i < NV && j < NV) return m2(i, j);
    if (i < NV) return -0.5 * mT(i, j - NV);
    if (j < NV) return -0.5 * mT(j, i - NV);
    return 0.25 * mQ(i - NV, j - NV);
  }
  // int phi_i dn phi_j
  __device__ __forceinline__ double N(int i, int j) const {
    if (i < NV && j < NV) return g(j) * m1(i);
    if (i < NV) {
      int a, b;
      edge_vertices<DIM>(j - NV, a, b);
      return -0.5 * (g(a) * m2(i, b) + g(b) * m2(i, a));
    }
    int a, b;
    edge_vertices<DIM>(i - NV, a, b);
    if (j < NV) return -0.5 * g(j) * m2(a, b);
    int c, d;
    edge_vertices<DIM>(j - NV, c, d);
    return 0.25 * (g(c) * mT(d, i - NV) + g(d) * mT(c, i - NV));
  }
  // int dn phi_i dn phi_j
  __device__ __forceinline__ double D(int i, int j) const {
    if (i < NV && j < NV) return g(i) * g(j) * m0();
    if (i < NV || j < NV) {
      const int v = i < NV ? i : j, ed = (i < NV ? j : i) - NV;
      int a, b;
      edge_vertices<DIM>(ed, a, b);
      return -0.5 * g(v) * (g(a) * m1(b) + g(b) * m1(a));
    }
    int a, b, c, d;
    edge_vertices<DIM>(i - NV, a, b);
    edge_vertices<DIM>(j - NV, c, d);
    return 0.25 * (g(a) * g(c) * m2(b, d) + g(a) * g(d) * m2(b, c) + g(b) * g(c) * m2(a, d) + g(b) * g(d) * m2(a, c));
  }
  // int grad phi_i . grad phi_j
  __device__ __forceinline__ double K(int i, int j) const {
    if (i < NV && j < NV) return gg(i, j) * m0();
    if (i < NV || j < NV) {
      const int v = i < NV ? i : j, ed = (i < NV ? j : i) - NV;
      int a, b;
      edge_vertices<DIM>(ed, a, b);
      return -0.5 * (gg(v, a) * m1(b) + gg(v, b) * m1(a));
    }
    int a, b, c, d;
    edge_vertices<DIM>(i - NV, a, b);
    edge_vertices<DIM>(j - NV, c, d);
    return 0.25 * (gg(a, c) * m2(b, d) + gg(a, d) * m2(b, c) + gg(b, c) * m2(a, d) + gg(b, d) * m2(a, c));
  }
};

template <int DIM, int ORDER>
__global__ void __launch_bounds__(XMT) k_xfem_mom(ngsx_mesh M, const double *__restrict__ lset, const CutTable *__restrict__ tab, int dt,
                                                 const double4 *__restrict__ rule_vol, int nr_vol, const double4 *__restrict__ rule_if, int nr_if,
                                                 XTermsDev X, int64_t nlist, const int32_t *__restrict__ els, const int8_t *__restrict__ xsign,
                                                 const double *__restrict__ ratio, int accumulate, double *__restrict__ mats,
                                                 const int32_t *__restrict__ rowslot) {
  using XM = XMom<DIM, ORDER>;
  constexpr int ND = XM::ND, NV = XM::NV, NEDGE = XM::NEDGE, N2C = 2 * ND;
  extern __shared__ __align__(16) unsigned char xsmem[];
  double4 *srule = reinterpret_cast<double4 *>(xsmem);                  // volume rule, then interface rule
  CutEntry *lut = reinterpret_cast<CutEntry *>(srule + nr_vol + nr_if);
  double *coefs = reinterpret_cast<double *>(lut + 16) + threadIdx.x * XCO;
  double *wmem = reinterpret_cast<double *>(lut + 16) + XMT * XCO + threadIdx.x * XM::TOT;
  for (int i = threadIdx.x; i < nr_vol; i += blockDim.x) srule[i] = rule_vol[i];
  for (int i = threadIdx.x; i < nr_if; i += blockDim.x) srule[nr_vol + i] = rule_if[i];
  if (threadIdx.x < 16) lut[threadIdx.x] = tab->e[DIM - 2][dt][threadIdx.x];
  __syncthreads();
  bool use0 = false, use1 = false, use2 = false, use3 = false, use4 = false;
  for (int t = 0; t < X.n; t++) {
    const int pr = X.pair[t];
    use0 |= pr == 0; use1 |= pr == 1; use2 |= pr == 2; use3 |= pr == 3; use4 |= pr == 4;
  }
  const bool IFACE = dt == IF;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nlist; k += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = els ? els[k] : k;
    double ls[4], p[DIM + 1][DIM];
    xload_el<DIM>(M, lset, e, ls, p);
    int mask;
    const int dom = sign_test<DIM>(ls, mask);
    const bool cut = dom == IF;
    double meas;
    {
      double m0 = 0.0, m1[NV], m2[XM::N2], mT[NV][XM::NT], mQ[XM::NQ];
#pragma unroll
      for (int a = 0; a < NV; a++) {
        m1[a] = 0.0;
#pragma unroll
        for (int t = 0; t < XM::NT; t++) mT[a][t] = 0.0;
      }
#pragma unroll
      for (int t = 0; t < XM::N2; t++) m2[t] = 0.0;
#pragma unroll
      for (int t = 0; t < XM::NQ; t++) mQ[t] = 0.0;
      ElGeom<DIM> G;
      element_geometry<DIM>(p, G);
      meas = fabs(G.det);
      double gl[DIM + 1][DIM];
#pragma unroll
      for (int d = 0; d < DIM; d++) {
        double s = 0.0;
#pragma unroll
        for (int kk = 0; kk < DIM; kk++) { gl[kk][d] = G.Jinv[kk][d]; s += G.Jinv[kk][d]; }
        gl[DIM][d] = -s;
      }
      double nrm[DIM], scale = 1.0;
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
        for (int kk = 0; kk < DIM; kk++) { const double v = ls[kk] - ls[DIM]; r2 += v * v; }
        const double inv = 1.0 / sqrt(g2);
#pragma unroll
        for (int d = 0; d < DIM; d++) nrm[d] *= inv;
        if (cut && IFACE) scale = sqrt(g2) / sqrt(r2);                  // |F^-T n_ref| (straightcutrule.cpp:558-576)
      }
      // geometry part of the slice: g_a = grad lam_a . n and the Gram matrix of the barycentric gradients
#pragma unroll
      for (int a = 0; a < NV; a++) {
        double v = 0.0;
#pragma unroll
        for (int d = 0; d < DIM; d++) v = fma(gl[a][d], nrm[d], v);
        wmem[XM::O_G + a] = v;
#pragma unroll
        for (int b = 0; b <= a; b++) {
          double w = 0.0;
#pragma unroll
          for (int d = 0; d < DIM; d++) w = fma(gl[a][d], gl[b][d], w);
          wmem[XM::O_GM + a * (a + 1) / 2 + b] = w;
        }
      }
      if (cut || dom == dt) {
        CutEntry E;
        if (cut) E = lut[mask]; else { E.ns = 1; E.npt = 0; }
        const double4 *rule = (cut && IFACE) ? srule + nr_vol : srule;
        const int nr = (cut && IFACE) ? nr_if : nr_vol;
        for (int sub = 0; sub < E.ns; sub++) {
          double P[4][3];
          double sfac = 1.0;
          if (cut) {
            if (IFACE) {
#pragma unroll
              for (int v = 0; v < DIM; v++) coded_point<DIM>(E.v[sub][v], ls, P[v]);
              sfac = simplex_factor(DIM, P);
            } else {
#pragma unroll
              for (int v = 0; v <= DIM; v++) coded_point<DIM>(E.v[sub][v], ls, P[v]);
              sfac = simplex_factor(DIM + 1, P);
            }
          }
          const double f0 = meas * sfac * scale;
          for (int q = 0; q < nr; q++) {
            const double4 r = rule[q];
            double xi[3] = {r.x, r.y, r.z};
            if (cut) { if (IFACE) simplex_point(DIM, P, r, xi); else simplex_point(DIM + 1, P, r, xi); }
            double lam[NV], fl[NV];
            double last = 1.0;
#pragma unroll
            for (int kk = 0; kk < DIM; kk++) { lam[kk] = xi[kk]; last -= xi[kk]; }
            lam[DIM] = last;
            const double f = f0 * r.w;
            m0 += f;
#pragma unroll
            for (int a = 0; a < NV; a++) {
              fl[a] = f * lam[a];
              m1[a] += fl[a];
#pragma unroll
              for (int b = 0; b <= a; b++) m2[a * (a + 1) / 2 + b] = fma(fl[a], lam[b], m2[a * (a + 1) / 2 + b]);
            }
            if constexpr (ORDER == 2) {
              double ep[XM::NT];
#pragma unroll
              for (int ed = 0; ed < NEDGE; ed++) {
                int a, b;
                edge_vertices<DIM>(ed, a, b);
                ep[ed] = lam[a] * lam[b];
              }
#pragma unroll
              for (int a = 0; a < NV; a++)
#pragma unroll
                for (int ed = 0; ed < NEDGE; ed++) mT[a][ed] = fma(fl[a], ep[ed], mT[a][ed]);
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
      wmem[0] = m0;
#pragma unroll
      for (int a = 0; a < NV; a++) {
        wmem[XM::O_M1 + a] = m1[a];
#pragma unroll
        for (int t = 0; t < XM::NT; t++) wmem[XM::O_MT + a * XM::NT + t] = mT[a][t];
      }
#pragma unroll
      for (int t = 0; t < XM::N2; t++) wmem[XM::O_M2 + t] = m2[t];
#pragma unroll
      for (int t = 0; t < XM::NQ; t++) wmem[XM::O_MQ + t] = mQ[t];
    }
    // the form's terms, folded per block class (test block * 3 + trial block) and operator pair
    {
      const double h = DIM == 2 ? sqrt(meas) : cbrt(meas);              // specialcf.mesh_size
      const double kap = __ldg(ratio + e);                              // CutRatioGF (xfem/cutinfo.cpp:91)
      for (int c = 0; c < 9; c++) {
        double acc[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
        for (int t = X.cls[c]; t < X.cls[c + 1]; t++) {
          const double kf = X.kappa[t] == 0 ? 1.0 : (X.kappa[t] == 1 ? kap : 1.0 - kap);
          const double v = X.coef[t] * xipow(h, X.hpow[t]) * kf;
          const int pr = X.pair[t];
#pragma unroll
          for (int q = 0; q < 5; q++) acc[q] += pr == q ? v : 0.0;
        }
#pragma unroll
        for (int q = 0; q < 5; q++) coefs[c * 5 + q] = acc[q];
      }
    }
    const XM W{wmem};
    const int8_t *sg = xsign + k * ND;
    double *out = mats + k * (N2C * N2C);
#pragma unroll 1
    for (int i = 0; i < ND; i++) {
      // packed rows (ngsx_set_packed_output): rows i and ND + i of block k go to the slots of their pairs (N2C is even)
      double *orow[2] = {out + i * N2C, out + (ND + i) * N2C};
      if (rowslot) {
        const int s0 = __ldg(rowslot + k * N2C + i), s1 = __ldg(rowslot + k * N2C + ND + i);
        orow[0] = s0 >= 0 ? mats + (int64_t)s0 * N2C : nullptr;
        orow[1] = s1 >= 0 ? mats + (int64_t)s1 * N2C : nullptr;
      }
      const int ci = 1 + (int)__ldg(sg + i);                            // block class of the x copy of dof i (1 + its sign)
#pragma unroll 1
      for (int j = 0; j < ND; j++) {
        const int cj = 1 + (int)__ldg(sg + j);
        // operator pairs (test, trial): 0 val-val, 1 val-dn, 2 dn-val, 3 dn-dn, 4 grad.grad
        const double b0 = use0 ? W.M(i, j) : 0.0, b1 = use1 ? W.N(i, j) : 0.0, b2 = use2 ? W.N(j, i) : 0.0;
        const double b3 = use3 ? W.D(i, j) : 0.0, b4 = use4 ? W.K(i, j) : 0.0;
#pragma unroll
        for (int bI = 0; bI < 2; bI++)
#pragma unroll
          for (int bJ = 0; bJ < 2; bJ++) {
            const int c = (bI ? ci : 0) * 3 + (bJ ? cj : 0);
            const double *cf = coefs + c * 5;
            double v = cf[0] * b0;
            v = fma(cf[1], b1, v); v = fma(cf[2], b2, v); v = fma(cf[3], b3, v); v = fma(cf[4], b4, v);
            if (orow[bI]) {
              double *o = orow[bI] + bJ * ND + j;
              *o = accumulate ? *o + v : v;
            }
          }
      }
    }
  }
}

template <int DIM, int ORDER>
static int launch_xfem(const ngsx_mesh &M, const double *lset, const DeviceData *dd, int dt, const double4 *rv, int nv, const double4 *ri, int ni,
                       const XTermsDev &X, int64_t nlist, const int32_t *els, const int8_t *xsign, const double *ratio, int accumulate, double *mats, const int32_t *rowslot,
                       cudaStream_t st) {
  const size_t smem = (size_t)(nv + ni) * sizeof(double4) + 16 * sizeof(CutEntry) + (size_t)XMT * (XCO + XMom<DIM, ORDER>::TOT) * sizeof(double);
  auto kern = k_xfem_mom<DIM, ORDER>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, XMT, smem));
  if (occ < 1) { set_error("xfem_elmat: kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
  const int64_t need = (nlist + XMT - 1) / XMT;
  const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)dd->sm_count * occ));
  kern<<<grid, XMT, smem, st>>>(M, lset, dd->cut_table, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, accumulate, mats, rowslot);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// DEFORMED mesh (row f-4): the same compound block with per-point geometry.  The monomial-moment formulation needs a constant
// Jacobian, so here a WARP takes an element: lane q evaluates rule point q of a round of XDQ (isoparametric map, shapes, mapped
// gradients, normal derivative, weight |det F| w [|F^-T n_ref|], the form's terms folded into 45 coefficients with the
// point's mesh_size) into the warp's shared-memory stage, then every lane accumulates its entries of the (2 nd)^2 block.
// Reference: DiffOpX on a deformed ElementTransformation (xfem/xfemdiffops.cpp:8-82 uses CalcMappedDShape(mip)).
// ---------------------------------------------------------------------------------------------------------------
constexpr int XDW = 4;            // warps per CTA
constexpr int XDQ = 16;           // rule points staged per round

struct XDefo { const int32_t *el2dof; const double *coef; };

template <int DIM, int ORDER, int DORD>
__global__ void __launch_bounds__(XDW * 32) k_xfem_def(ngsx_mesh M, XDefo D, const double *__restrict__ lset, const CutTable *__restrict__ tab, int dt,
                                                      const double4 *__restrict__ rule_vol, int nr_vol, const double4 *__restrict__ rule_if,
                                                      int nr_if, XTermsDev X, int64_t nlist, const int32_t *__restrict__ els,
                                                      const int8_t *__restrict__ xsign, const double *__restrict__ ratio, int accumulate,
                                                      double *__restrict__ mats, const int32_t *__restrict__ rowslot) {
  constexpr int ND = H1<DIM, ORDER>::ND, N2C = 2 * ND, NE = N2C * N2C, NDG = H1<DIM, DORD>::ND, T = (NE + 31) / 32;
  constexpr size_t PER_WARP = (size_t)XDQ * ND * 32 + ((size_t)XDQ * ND + XDQ + XDQ * XCO + NDG * DIM) * 8;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char *wbase = smem_raw + warp * PER_WARP;
  double4 *stage = reinterpret_cast<double4 *>(wbase);
  double *dnq = reinterpret_cast<double *>(wbase + (size_t)XDQ * ND * 32);
  double *facq = dnq + XDQ * ND;
  double *cq = facq + XDQ;
  double *gc = cq + XDQ * XCO;
  const bool IFACE = dt == IF;
  const int64_t nwarps = (int64_t)gridDim.x * XDW;
  for (int64_t k = (int64_t)blockIdx.x * XDW + warp; k < nlist; k += nwarps) {
    const int64_t e = els ? els[k] : k;
    const int32_t *ev = M.el2v + e * (DIM + 1);
    double ls[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int a = 0; a <= DIM; a++) ls[a] = __ldg(lset + __ldg(ev + a));
    __syncwarp();
    for (int t = lane; t < NDG * DIM; t += 32) {
      const int i = t / DIM, d = t % DIM;
      double c = __ldg(D.coef + (int64_t)__ldg(D.el2dof + e * NDG + i) * DIM + d);
      if (i <= DIM) c += __ldg(M.vert + (int64_t)__ldg(ev + i) * DIM + d);
      gc[t] = c;
    }
    __syncwarp();
    int mask;
    const int dom = sign_test<DIM>(ls, mask);
    const bool cut = dom == IF;
    // this lane's entries (I, J) of the compound block and their block classes
    const int8_t *sg = xsign + k * ND;
    int ei[T], ej[T], ec[T];
#pragma unroll
    for (int t = 0; t < T; t++) {
      const int idx = lane + 32 * t;
      const int I = idx < NE ? idx / N2C : 0, J = idx < NE ? idx % N2C : 0;
      const int bI = I >= ND, bJ = J >= ND;
      ei[t] = I - bI * ND; ej[t] = J - bJ * ND;
      ec[t] = (bI ? 1 + (int)__ldg(sg + ei[t]) : 0) * 3 + (bJ ? 1 + (int)__ldg(sg + ej[t]) : 0);
    }
    double acc[T];
#pragma unroll
    for (int t = 0; t < T; t++) acc[t] = 0.0;
    if (cut || dom == dt) {
      const double kap = __ldg(ratio + e);
      double r2 = 0.0;
#pragma unroll
      for (int a = 0; a < DIM; a++) { const double v = ls[a] - ls[DIM]; r2 += v * v; }
      const bool iface = cut && IFACE;
      CutEntry E;
      if (cut) E = tab->e[DIM - 2][dt][mask]; else { E.ns = 1; E.npt = 0; }
      const double4 *rule = iface ? rule_if : rule_vol;
      const int nr = iface ? nr_if : nr_vol;
      for (int sidx = 0; sidx < E.ns; sidx++) {
        double P[4][3];
        double sfac = 1.0;
        if (cut) {
          for (int v = 0; v < E.npt; v++) coded_point<DIM>(E.v[sidx][v], ls, P[v]);
          sfac = simplex_factor(E.npt, P);
        }
        for (int q0 = 0; q0 < nr; q0 += XDQ) {
          const int nq = min(XDQ, nr - q0);
          if (lane < nq) {
            const double4 r = rule[q0 + lane];
            double xi[3] = {r.x, r.y, r.z};
            if (cut) simplex_point(E.npt, P, r, xi);
            double lam[4] = {xi[0], xi[1], xi[2], 0.0}, last = 1.0;
#pragma unroll
            for (int a = 0; a < DIM; a++) last -= lam[a];
            lam[DIM] = last;
            double x[DIM], gl[DIM + 1][DIM], det;
            iso_point<DIM, DORD>(GcWarp{gc}, lam, x, gl, det);
            const double adet = fabs(det);
            double nrm[DIM], g2 = 0.0;
#pragma unroll
            for (int d = 0; d < DIM; d++) {
              double v = 0.0;
#pragma unroll
              for (int a = 0; a <= DIM; a++) v += ls[a] * gl[a][d];
              nrm[d] = v; g2 += v * v;
            }
            const double gn = sqrt(g2), inv = 1.0 / gn;
#pragma unroll
            for (int d = 0; d < DIM; d++) nrm[d] *= inv;
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
            facq[lane] = adet * (r.w * sfac * (iface ? gn / sqrt(r2) : 1.0));
            const double h = DIM == 2 ? sqrt(adet) : cbrt(adet);        // specialcf.mesh_size at this point
            for (int c = 0; c < 9; c++) {
              double a5[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
              for (int t = X.cls[c]; t < X.cls[c + 1]; t++) {
                const double kf = X.kappa[t] == 0 ? 1.0 : (X.kappa[t] == 1 ? kap : 1.0 - kap);
                const double v = X.coef[t] * xipow(h, X.hpow[t]) * kf;
                const int pr = X.pair[t];
#pragma unroll
                for (int qq = 0; qq < 5; qq++) a5[qq] += pr == qq ? v : 0.0;
              }
#pragma unroll
              for (int qq = 0; qq < 5; qq++) cq[lane * XCO + c * 5 + qq] = a5[qq];
            }
          }
          __syncwarp();
          for (int q = 0; q < nq; q++) {
            const double f = facq[q];
#pragma unroll
            for (int t = 0; t < T; t++) {
              const double4 bi = stage[q * ND + ei[t]], bj = stage[q * ND + ej[t]];
              const double di = dnq[q * ND + ei[t]], dj = dnq[q * ND + ej[t]];
              const double *cf = cq + q * XCO + ec[t] * 5;
              const double gg = bi.y * bj.y + bi.z * bj.z + (DIM == 3 ? bi.w * bj.w : 0.0);
              // operator pairs (test, trial): 0 val-val, 1 val-dn, 2 dn-val, 3 dn-dn, 4 grad.grad
              double v = cf[0] * (bi.x * bj.x);
              v = fma(cf[1], bi.x * dj, v); v = fma(cf[2], di * bj.x, v); v = fma(cf[3], di * dj, v); v = fma(cf[4], gg, v);
              acc[t] += f * v;
            }
          }
          __syncwarp();
        }
      }
    }
#pragma unroll
    for (int t = 0; t < T; t++) {
      const int idx = lane + 32 * t;
      if (idx < NE) {
        const int I = idx / N2C, J = idx % N2C;
        double *o = mats + k * NE + idx;
        if (rowslot) {
          const int sl = __ldg(rowslot + k * N2C + I);
          o = sl >= 0 ? mats + (int64_t)sl * N2C + J : nullptr;
        }
        if (o) *o = accumulate ? *o + acc[t] : acc[t];
      }
    }
  }
}

template <int DIM, int ORDER, int DORD>
static int launch_xfem_def(const ngsx_mesh &M, const XDefo &D, const double *lset, const DeviceData *dd, int dt, const double4 *rv, int nv,
                           const double4 *ri, int ni, const XTermsDev &X, int64_t nlist, const int32_t *els, const int8_t *xsign,
                           const double *ratio, int accumulate, double *mats, const int32_t *rowslot, cudaStream_t st) {
  constexpr int ND = H1<DIM, ORDER>::ND, NDG = H1<DIM, DORD>::ND;
  const size_t smem = XDW * ((size_t)XDQ * ND * 32 + ((size_t)XDQ * ND + XDQ + XDQ * XCO + NDG * DIM) * 8);
  auto kern = k_xfem_def<DIM, ORDER, DORD>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t need = (nlist + XDW - 1) / XDW;
  const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)dd->sm_count * 8));
  kern<<<grid, XDW * 32, smem, st>>>(M, D, lset, dd->cut_table, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, accumulate, mats, rowslot);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

}  // namespace ngsx

using namespace ngsx;

static int prepare_xterms(int32_t nterms, const ngsx_xterm *h_terms, XTermsDev &X) {
  X.n = nterms;
  for (int t = 0; t < nterms; t++) {
    const ngsx_xterm &s = h_terms[t];
    NGSX_REQUIRE(s.pair >= 0 && s.pair <= 4 && s.kappa >= 0 && s.kappa <= 2 && s.blk_u >= 0 && s.blk_u <= 2 && s.blk_v >= 0 && s.blk_v <= 2 &&
                 s.hpow >= -8 && s.hpow <= 8, "xfem_elmat: bad term");
  }
  {
    // stable counting sort by (test block, trial block): an entry of the compound block only walks its own class; the order of
    // the terms inside a class -- the summation order -- is the caller's
    int o = 0;
    for (int c = 0; c < 9; c++) {
      X.cls[c] = (int8_t)o;
      for (int t = 0; t < nterms; t++) {
        const ngsx_xterm &s = h_terms[t];
        if (s.blk_v * 3 + s.blk_u != c) continue;
        X.coef[o] = s.coef; X.hpow[o] = (int8_t)s.hpow; X.kappa[o] = (int8_t)s.kappa; X.pair[o] = (int8_t)s.pair;
        X.blk_u[o] = (int8_t)s.blk_u; X.blk_v[o] = (int8_t)s.blk_v;
        o++;
      }
    }
    X.cls[9] = (int8_t)o;
  }
  return NGSX_OK;
}

extern "C" int ngsx_xfem_elmat_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, int32_t nterms,
                                  const ngsx_xterm *h_terms, int64_t nlist, const int32_t *els, const int8_t *xsign, const double *ratio,
                                  int accumulate, double *mats, void *stream) {
  NGSX_REQUIRE(mesh && (mesh->dim == 2 || mesh->dim == 3) && mesh->vert && mesh->el2v && lset_v, "xfem_elmat: bad mesh / level set");
  NGSX_REQUIRE(dt >= 0 && dt <= 2 && nlist >= 0 && nterms >= 0 && (nterms == 0 || h_terms), "xfem_elmat: bad arguments");
  NGSX_REQUIRE(nlist == 0 || (xsign && ratio && mats), "xfem_elmat: NULL array");
  if (nterms > XMAXTERMS) { set_error("xfem_elmat: more than 32 product terms"); return NGSX_ERR_UNSUPPORTED; }
  if (fe_order < 1 || fe_order > 2) { set_error("xfem_elmat: H1 order outside the CUDA path (1, 2); no CPU fallback"); return NGSX_ERR_UNSUPPORTED; }
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  const double4 *rv, *ri; int nv, ni;
  if (!rule_lookup(dd, mesh->dim, intorder, &rv, &nv) || !rule_lookup(dd, mesh->dim - 1, intorder, &ri, &ni)) {
    set_error("xfem_elmat: integration order beyond the tabulated reference rules");
    return NGSX_ERR_UNSUPPORTED;
  }
  XTermsDev X{};
  if (int rc = prepare_xterms(nterms, h_terms, X)) return rc;
  const int32_t *rowslot = take_packed_output();
  if (nlist == 0) return NGSX_OK;
  cudaStream_t st = as_stream(stream);
  switch (mesh->dim * 10 + fe_order) {
    case 21: return launch_xfem<2, 1>(*mesh, lset_v, dd, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, accumulate, mats, rowslot, st);
    case 22: return launch_xfem<2, 2>(*mesh, lset_v, dd, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, accumulate, mats, rowslot, st);
    case 31: return launch_xfem<3, 1>(*mesh, lset_v, dd, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, accumulate, mats, rowslot, st);
    case 32: return launch_xfem<3, 2>(*mesh, lset_v, dd, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, accumulate, mats, rowslot, st);
  }
  return NGSX_ERR_UNSUPPORTED;
}

extern "C" int ngsx_xfem_elmat_h1_deformed(const ngsx_mesh *mesh, const ngsx_deformation *defo, const double *lset_v, int fe_order, int dt,
                                           int intorder, int32_t nterms, const ngsx_xterm *h_terms, int64_t nlist, const int32_t *els,
                                           const int8_t *xsign, const double *ratio, int accumulate, double *mats, void *stream) {
  NGSX_REQUIRE(mesh && (mesh->dim == 2 || mesh->dim == 3) && mesh->vert && mesh->el2v && lset_v, "xfem_elmat_deformed: bad mesh / level set");
  NGSX_REQUIRE(defo && defo->el2dof && defo->coef, "xfem_elmat_deformed: bad deformation");
  NGSX_REQUIRE(dt >= 0 && dt <= 2 && nlist >= 0 && nterms >= 0 && (nterms == 0 || h_terms), "xfem_elmat_deformed: bad arguments");
  NGSX_REQUIRE(nlist == 0 || (xsign && ratio && mats), "xfem_elmat_deformed: NULL array");
  if (nterms > XMAXTERMS) { set_error("xfem_elmat_deformed: more than 32 product terms"); return NGSX_ERR_UNSUPPORTED; }
  if (fe_order < 1 || fe_order > 2 || defo->order < 1 || defo->order > 2) {
    set_error("xfem_elmat_deformed: H1 / deformation order outside the CUDA path (1, 2); no CPU fallback");
    return NGSX_ERR_UNSUPPORTED;
  }
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  const double4 *rv, *ri; int nv, ni;
  if (!rule_lookup(dd, mesh->dim, intorder, &rv, &nv) || !rule_lookup(dd, mesh->dim - 1, intorder, &ri, &ni)) {
    set_error("xfem_elmat_deformed: integration order beyond the tabulated reference rules");
    return NGSX_ERR_UNSUPPORTED;
  }
  XTermsDev X{};
  if (int rc = prepare_xterms(nterms, h_terms, X)) return rc;
  const int32_t *rowslot = take_packed_output();
  if (nlist == 0) return NGSX_OK;
  cudaStream_t st = as_stream(stream);
  const XDefo D{defo->el2dof, defo->coef};
#define XDEF(DIM, ORD, DORD) launch_xfem_def<DIM, ORD, DORD>(*mesh, D, lset_v, dd, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, accumulate, mats, rowslot, st)
  switch (mesh->dim * 100 + fe_order * 10 + defo->order) {
    case 211: return XDEF(2, 1, 1);
    case 212: return XDEF(2, 1, 2);
    case 221: return XDEF(2, 2, 1);
    case 222: return XDEF(2, 2, 2);
    case 311: return XDEF(3, 1, 1);
    case 312: return XDEF(3, 1, 2);
    case 321: return XDEF(3, 2, 1);
    case 322: return XDEF(3, 2, 2);
  }
#undef XDEF
  return NGSX_ERR_UNSUPPORTED;
}
