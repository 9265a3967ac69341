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
// One warp per listed element: lane q evaluates quadrature point q of a round into shared memory, the nd^2 entries of each
// base matrix are spread over the lanes; the compound block leaves as coalesced stores.  The blocks are scattered by the
// owner-computes gather (ngsx_assemble_csr_rows_mats).
#include <algorithm>
#include "ngsx_host.cuh"

namespace ngsx {

constexpr int XW = 4;             // warps per CTA
constexpr int XQ = 16;            // quadrature points staged per round
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

template <int DIM, int ORDER>
__global__ void __launch_bounds__(XW * 32) k_xfem(ngsx_mesh M, const double *__restrict__ lset, const CutTable *__restrict__ tab, int dt,
                                                  const double4 *__restrict__ rule_vol, int nr_vol, const double4 *__restrict__ rule_if, int nr_if,
                                                  XTermsDev X, int64_t nlist, const int32_t *__restrict__ els, const int8_t *__restrict__ xsign,
                                                  const double *__restrict__ ratio, double *__restrict__ mats) {
  constexpr int ND = H1<DIM, ORDER>::ND, NE = ND * ND, N2 = 2 * ND, T = (NE + 31) / 32;
  extern __shared__ __align__(16) unsigned char xsmem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // per warp: stage [XQ][ND] x (value, dn, gx, gy, gz) + weights [XQ] + base matrices [4][NE] + term coefficients
  double *wbase = reinterpret_cast<double *>(xsmem) + warp * (XQ * ND * 5 + XQ + 4 * NE + XMAXTERMS);
  double *stage = wbase, *facq = wbase + XQ * ND * 5, *base = facq + XQ, *tc = base + 4 * NE;
  const bool need_k = [&] { bool b = false; for (int t = 0; t < X.n; t++) b |= X.pair[t] == 4; return b; }();
  const bool need_b = [&] { bool b = false; for (int t = 0; t < X.n; t++) b |= X.pair[t] != 4; return b; }();
  const int64_t nwarps = (int64_t)gridDim.x * XW;
  for (int64_t k = (int64_t)blockIdx.x * XW + warp; k < nlist; k += nwarps) {
    const int64_t e = els ? els[k] : k;
    double ls[4], p[DIM + 1][DIM];
    xload_el<DIM>(M, lset, e, ls, p);
    int mask;
    const int dom = sign_test<DIM>(ls, mask);
    const bool cut = dom == IF;
    double aM[T], aN[T], aD[T], aK[T];
#pragma unroll
    for (int t = 0; t < T; t++) aM[t] = aN[t] = aD[t] = aK[t] = 0.0;
    ElGeom<DIM> G;
    element_geometry<DIM>(p, G);
    const double meas = fabs(G.det);
    if (cut || dom == dt) {
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
        if (cut && dt == IF) scale = sqrt(g2) / sqrt(r2);               // |F^-T n_ref| (straightcutrule.cpp:558-576)
      }
      CutEntry E;
      if (cut) E = tab->e[DIM - 2][dt][mask]; else { E.ns = 1; E.npt = 0; }
      const double4 *rule = (cut && dt == IF) ? rule_if : rule_vol;
      const int nr = (cut && dt == IF) ? nr_if : nr_vol;
      for (int s = 0; s < E.ns; s++) {
        double P[4][3];
        double sfac = 1.0;
        if (cut) {
          for (int v = 0; v < E.npt; v++) coded_point<DIM>(E.v[s][v], ls, P[v]);
          sfac = simplex_factor(E.npt, P);
        }
        for (int q0 = 0; q0 < nr; q0 += XQ) {
          const int nq = min(XQ, nr - q0);
          if (lane < nq) {
            const double4 r = rule[q0 + lane];
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
              h1_shape_grad<DIM, ORDER>(i, lam, gl, sv, gv);
              double dn = 0.0;
#pragma unroll
              for (int d = 0; d < DIM; d++) dn += gv[d] * nrm[d];
              double *o = stage + (lane * ND + i) * 5;
              o[0] = sv; o[1] = dn; o[2] = gv[0]; o[3] = gv[1]; o[4] = DIM == 3 ? gv[DIM - 1] : 0.0;
            }
            facq[lane] = meas * (r.w * sfac * scale);
          }
          __syncwarp();
          for (int q = 0; q < nq; q++) {
            const double f = facq[q];
#pragma unroll
            for (int t = 0; t < T; t++) {
              const int idx = lane + 32 * t;
              if (idx < NE) {
                const double *bi = stage + (q * ND + idx / ND) * 5, *bj = stage + (q * ND + idx % ND) * 5;
                if (need_b) {
                  aM[t] += f * bi[0] * bj[0];
                  aN[t] += f * bi[0] * bj[1];
                  aD[t] += f * bi[1] * bj[1];
                }
                if (need_k) aK[t] += f * (bi[2] * bj[2] + bi[3] * bj[3] + bi[4] * bj[4]);
              }
            }
          }
          __syncwarp();
        }
      }
    }
#pragma unroll
    for (int t = 0; t < T; t++) {
      const int idx = lane + 32 * t;
      if (idx < NE) { base[idx] = aM[t]; base[NE + idx] = aN[t]; base[2 * NE + idx] = aD[t]; base[3 * NE + idx] = aK[t]; }
    }
    if (lane < X.n) {
      const double h = DIM == 2 ? sqrt(meas) : cbrt(meas);                // specialcf.mesh_size
      const double kap = __ldg(ratio + e);                                // CutRatioGF (xfem/cutinfo.cpp:91)
      const double kf = X.kappa[lane] == 0 ? 1.0 : (X.kappa[lane] == 1 ? kap : 1.0 - kap);
      tc[lane] = X.coef[lane] * xipow(h, X.hpow[lane]) * kf;
    }
    __syncwarp();
    // compound block: row I = test function, column J = trial function; index = block * ND + local dof
    const int8_t *sg = xsign + k * ND;
    double *out = mats + k * (N2 * N2);
    for (int idx = lane; idx < N2 * N2; idx += 32) {
      const int I = idx / N2, J = idx - I * N2;
      const int bI = I >= ND, i = I - bI * ND, bJ = J >= ND, j = J - bJ * ND;
      // blk 0: std dofs; blk 1 / 2: x dofs whose sign is NEG / POS
      const int cI = bI ? 1 + (int)sg[i] : 0, cJ = bJ ? 1 + (int)sg[j] : 0;
      double v = 0.0;
      const int c = cI * 3 + cJ;
      for (int t = X.cls[c]; t < X.cls[c + 1]; t++) {
        const int pr = X.pair[t];
        const double b = pr == 0 ? base[i * ND + j] : (pr == 1 ? base[NE + i * ND + j] : (pr == 2 ? base[NE + j * ND + i]
                         : (pr == 3 ? base[2 * NE + i * ND + j] : base[3 * NE + i * ND + j])));
        v += tc[t] * b;
      }
      out[idx] = v;
    }
    __syncwarp();
  }
}

template <int DIM, int ORDER>
static int launch_xfem(const ngsx_mesh &M, const double *lset, const DeviceData *dd, int dt, const double4 *rv, int nv, const double4 *ri, int ni,
                       const XTermsDev &X, int64_t nlist, const int32_t *els, const int8_t *xsign, const double *ratio, double *mats, cudaStream_t st) {
  constexpr int ND = H1<DIM, ORDER>::ND, NE = ND * ND;
  const size_t smem = (size_t)XW * (XQ * ND * 5 + XQ + 4 * NE + XMAXTERMS) * sizeof(double);
  auto kern = k_xfem<DIM, ORDER>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t need = (nlist + XW - 1) / XW;
  const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)dd->sm_count * 8));
  kern<<<grid, XW * 32, smem, st>>>(M, lset, dd->cut_table, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, mats);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

}  // namespace ngsx

using namespace ngsx;

extern "C" int ngsx_xfem_elmat_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, int32_t nterms,
                                  const ngsx_xterm *h_terms, int64_t nlist, const int32_t *els, const int8_t *xsign, const double *ratio,
                                  double *mats, void *stream) {
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
  if (nlist == 0) return NGSX_OK;
  cudaStream_t st = as_stream(stream);
  switch (mesh->dim * 10 + fe_order) {
    case 21: return launch_xfem<2, 1>(*mesh, lset_v, dd, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, mats, st);
    case 22: return launch_xfem<2, 2>(*mesh, lset_v, dd, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, mats, st);
    case 31: return launch_xfem<3, 1>(*mesh, lset_v, dd, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, mats, st);
    case 32: return launch_xfem<3, 2>(*mesh, lset_v, dd, dt, rv, nv, ri, ni, X, nlist, els, xsign, ratio, mats, st);
  }
  return NGSX_ERR_UNSUPPORTED;
}
