// ngsx_facetpatch.cu -- ghost-penalty facet-patch matrices (c-5):
//   coef * h^hpow * (u - u.Other()) (v - v.Other())  integrated over BOTH elements of a facet patch
// (SymbolicFacetPatchBilinearFormIntegrator::CalcFacetMatrix, xfem/symboliccutbfi.cpp:1202-1411; the tutorial form
// py_tutorials/fictdom_ghostpen.py:105-107).  Volume rule of order 2*fe_order on each element (:1219-1220); every point
// is evaluated in both elements -- the Newton pull-back of MapPatchIntegrationPoint (:1115-1199) is one exact
// affine solve here.  One warp per patch: lane q stages the "jump vector" of point q in shared memory, then the
// (2 nd)^2 entries are distributed over the lanes.
#include "ngsx_host.cuh"
#include "ngsx_iso.cuh"

namespace ngsx {

constexpr int FP_WARPS = 4;

template <int DIM>
__device__ __forceinline__ void load_el_geo(const ngsx_mesh &M, int64_t e, double (*p)[DIM]) {
  for (int k = 0; k <= DIM; k++) {
    int v = __ldg(M.el2v + e * (DIM + 1) + k);
#pragma unroll
    for (int d = 0; d < DIM; d++) p[k][d] = __ldg(M.vert + (int64_t)v * DIM + d);
  }
}

template <int DIM, int ORDER>
__global__ void __launch_bounds__(FP_WARPS * 32) k_facetpatch(ngsx_mesh M, const double4 *__restrict__ rule, int nr, double coef, int hpow,
                                                             int64_t npatch, const int32_t *__restrict__ patch_els, double *__restrict__ elmats) {
  constexpr int ND = H1<DIM, ORDER>::ND, N2 = 2 * ND, NE = N2 * N2, T = (NE + 31) / 32;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double4 *srule = reinterpret_cast<double4 *>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *stage = reinterpret_cast<double *>(smem_raw + (size_t)nr * 32) + warp * (32 * N2 + 32);
  double *facq = stage + 32 * N2;
  for (int i = threadIdx.x; i < nr; i += blockDim.x) srule[i] = rule[i];
  __syncthreads();
  const int64_t nwarps = (int64_t)gridDim.x * FP_WARPS;
  for (int64_t k = (int64_t)blockIdx.x * FP_WARPS + warp; k < npatch; k += nwarps) {
    double p[2][DIM + 1][DIM];
    ElGeom<DIM> G[2];
    double hfac[2];
    for (int s = 0; s < 2; s++) {
      load_el_geo<DIM>(M, patch_els[2 * k + s], p[s]);
      element_geometry<DIM>(p[s], G[s]);
    }
    {
      // the coefficient is evaluated on mir1 for ALL patch points (symboliccutbfi.cpp:1359): h = mesh_size of element 0;
      // the weight is the measure of the element the point lives on (mir1.GetWeight() after MapPatchIntegrationPoint)
      double m0 = fabs(G[0].det);
      double h = DIM == 2 ? sqrt(m0) : cbrt(m0);
      double c = coef * ipow(h, hpow);
      hfac[0] = c * m0;
      hfac[1] = c * fabs(G[1].det);
    }
    double acc[T];
#pragma unroll
    for (int t = 0; t < T; t++) acc[t] = 0.0;
    const int ntot = 2 * nr;
    for (int q0 = 0; q0 < ntot; q0 += 32) {
      const int nq = min(32, ntot - q0);
      if (lane < nq) {
        const int q = q0 + lane, side = q >= nr ? 1 : 0;
        const double4 r = srule[q - side * nr];
        double lam[4] = {r.x, r.y, r.z, 0.0}, last = 1.0;
#pragma unroll
        for (int kk = 0; kk < DIM; kk++) last -= lam[kk];
        lam[DIM] = last;
        double x[DIM];
#pragma unroll
        for (int d = 0; d < DIM; d++) {
          double v = 0.0;
#pragma unroll
          for (int kk = 0; kk <= DIM; kk++) v += lam[kk] * p[side][kk][d];
          x[d] = v;
        }
        // barycentrics of x in the other element: xi = Jinv_o (x - p_o[DIM])
        double lo[4] = {0.0, 0.0, 0.0, 0.0}, lasto = 1.0;
#pragma unroll
        for (int kk = 0; kk < DIM; kk++) {
          double v = 0.0;
#pragma unroll
          for (int d = 0; d < DIM; d++) v += G[1 - side].Jinv[kk][d] * (x[d] - p[1 - side][DIM][d]);
          lo[kk] = v; lasto -= v;
        }
        lo[DIM] = lasto;
        for (int i = 0; i < ND; i++) {
          stage[lane * N2 + side * ND + i] = h1_shape<DIM, ORDER>(i, lam);
          stage[lane * N2 + (1 - side) * ND + i] = -h1_shape<DIM, ORDER>(i, lo);
        }
        facq[lane] = hfac[side] * r.w;
      }
      __syncwarp();
      for (int q = 0; q < nq; q++) {
        const double f = facq[q];
#pragma unroll
        for (int t = 0; t < T; t++) {
          int idx = lane + 32 * t;
          if (idx < NE) acc[t] += f * stage[q * N2 + idx / N2] * stage[q * N2 + idx % N2];
        }
      }
      __syncwarp();
    }
#pragma unroll
    for (int t = 0; t < T; t++) {
      int idx = lane + 32 * t;
      if (idx < NE) elmats[k * NE + idx] = acc[t];
    }
  }
}

template <int DIM, int ORDER>
static int launch_fp(const ngsx_mesh &M, const DeviceData *dd, const double4 *rule, int nr, double coef, int hpow, int64_t np,
                     const int32_t *patch_els, double *elmats, cudaStream_t st) {
  constexpr int N2 = 2 * H1<DIM, ORDER>::ND;
  size_t smem = (size_t)nr * 32 + (size_t)FP_WARPS * (32 * N2 + 32) * 8;
  auto kern = k_facetpatch<DIM, ORDER>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t need = (np + FP_WARPS - 1) / FP_WARPS;
  unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)dd->sm_count * 8));
  kern<<<grid, FP_WARPS * 32, smem, st>>>(M, rule, nr, coef, hpow, np, patch_els, elmats);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

// ---- merged ("compressed") patch blocks: the form the row gather consumes --------------------------------------------
// NGSolve merges the dofs the two elements of a patch share before AddElementMatrix (compressed_dnums / compressed_elmat,
// NGS/comp/bilinearform.cpp:1800-1853).  Here the merge happens before the quadrature: cidx[a] is the position of local
// dof a of [dofs(el0) | dofs(el1)] in the merged list (dofs of el0, then the dofs of el1 not in el0), and the kernel
// integrates the merged jump vector directly -- (2 nd - shared)^2 instead of (2 nd)^2 products per point.
__global__ void __launch_bounds__(128) k_fp_table(int64_t npatch, const int32_t *__restrict__ patch_els, int nd, const int32_t *__restrict__ el2dof,
                                                 int ndp, int32_t *__restrict__ table, uint8_t *__restrict__ cidx, int *__restrict__ bad) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= npatch) return;
  const int32_t *d0 = el2dof + (int64_t)patch_els[2 * k] * nd, *d1 = el2dof + (int64_t)patch_els[2 * k + 1] * nd;
  int32_t *t = table + k * ndp;
  uint8_t *c = cidx + k * 2 * nd;
  for (int a = 0; a < nd; a++) { if (a < ndp) t[a] = d0[a]; c[a] = (uint8_t)a; }
  int n = nd;
  for (int b = 0; b < nd; b++) {
    const int32_t dof = d1[b];
    int hit = -1;
    for (int a = 0; a < nd; a++) if (d0[a] == dof) hit = a;
    if (hit < 0) { hit = n; if (n < ndp) t[n] = dof; n++; }
    c[nd + b] = (uint8_t)hit;
  }
  if (n != ndp) atomicOr(bad, 1);
}

template <int DIM, int ORDER>
__global__ void __launch_bounds__(FP_WARPS * 32) k_facetpatch_c(ngsx_mesh M, const double4 *__restrict__ rule, int nr, double coef, int hpow,
                                                               int64_t npatch, const int32_t *__restrict__ patch_els,
                                                               const uint8_t *__restrict__ cidx, double *__restrict__ mats,
                                                               const int32_t *__restrict__ rowslot) {
  // dofs on the shared facet: vertices + (order - 1) per edge (+ one face function of a tet face at order 3)
  constexpr int ND = H1<DIM, ORDER>::ND, NFD = DIM == 2 ? ORDER + 1 : (ORDER == 1 ? 3 : (ORDER == 2 ? 6 : 10));
  constexpr int NP = 2 * ND - NFD, NE = NP * NP, NS = NP * (NP + 1) / 2, T = (NS + 31) / 32;   // symmetric: upper triangle only
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double4 *srule = reinterpret_cast<double4 *>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *stage = reinterpret_cast<double *>(smem_raw + (size_t)nr * 32) + warp * (32 * NP + 32);
  double *facq = stage + 32 * NP;
  for (int i = threadIdx.x; i < nr; i += blockDim.x) srule[i] = rule[i];
  __syncthreads();
  int ei[T], ej[T];                                  // this lane's upper-triangle entries
#pragma unroll
  for (int t = 0; t < T; t++) {
    int idx = lane + 32 * t, i = 0;
    if (idx >= NS) idx = 0;
    while (idx >= NP - i) { idx -= NP - i; i++; }
    ei[t] = i; ej[t] = i + idx;
  }
  const int64_t nwarps = (int64_t)gridDim.x * FP_WARPS;
  for (int64_t k = (int64_t)blockIdx.x * FP_WARPS + warp; k < npatch; k += nwarps) {
    double p[2][DIM + 1][DIM];
    ElGeom<DIM> G[2];
    double hfac[2];
    unsigned orient[2] = {0u, 0u};                  // order 3: edge orientations from the global vertex numbers of each element
    for (int s = 0; s < 2; s++) {
      load_el_geo<DIM>(M, patch_els[2 * k + s], p[s]);
      element_geometry<DIM>(p[s], G[s]);
      if constexpr (ORDER == 3) {
        int vn[4] = {0, 0, 0, 0};
        for (int m = 0; m <= DIM; m++) vn[m] = __ldg(M.el2v + (int64_t)patch_els[2 * k + s] * (DIM + 1) + m);
        orient[s] = edge_orientation<DIM>(vn);
      }
    }
    {
      const double m0 = fabs(G[0].det);
      const double h = DIM == 2 ? sqrt(m0) : cbrt(m0);
      const double c = coef * ipow(h, hpow);
      hfac[0] = c * m0;
      hfac[1] = c * fabs(G[1].det);
    }
    const uint8_t *ci = cidx + k * 2 * ND;
    double acc[T];
#pragma unroll
    for (int t = 0; t < T; t++) acc[t] = 0.0;
    const int ntot = 2 * nr;
    for (int q0 = 0; q0 < ntot; q0 += 32) {
      const int nq = min(32, ntot - q0);
      if (lane < nq) {
        const int q = q0 + lane, side = q >= nr ? 1 : 0;
        const double4 r = srule[q - side * nr];
        double lam[4] = {r.x, r.y, r.z, 0.0}, last = 1.0;
#pragma unroll
        for (int kk = 0; kk < DIM; kk++) last -= lam[kk];
        lam[DIM] = last;
        double x[DIM];
#pragma unroll
        for (int d = 0; d < DIM; d++) {
          double v = 0.0;
#pragma unroll
          for (int kk = 0; kk <= DIM; kk++) v += lam[kk] * p[side][kk][d];
          x[d] = v;
        }
        double lo[4] = {0.0, 0.0, 0.0, 0.0}, lasto = 1.0;
#pragma unroll
        for (int kk = 0; kk < DIM; kk++) {
          double v = 0.0;
#pragma unroll
          for (int d = 0; d < DIM; d++) v += G[1 - side].Jinv[kk][d] * (x[d] - p[1 - side][DIM][d]);
          lo[kk] = v; lasto -= v;
        }
        lo[DIM] = lasto;
        double *jc = stage + lane * NP;
        for (int c = 0; c < NP; c++) jc[c] = 0.0;
        for (int i = 0; i < ND; i++) {
          jc[ci[side * ND + i]] += h1_shape<DIM, ORDER>(i, lam, orient[side]);
          jc[ci[(1 - side) * ND + i]] -= h1_shape<DIM, ORDER>(i, lo, orient[1 - side]);
        }
        facq[lane] = hfac[side] * r.w;
      }
      __syncwarp();
      for (int q = 0; q < nq; q++) {
        const double f = facq[q];
#pragma unroll
        for (int t = 0; t < T; t++) {
          acc[t] += f * stage[q * NP + ei[t]] * stage[q * NP + ej[t]];
        }
      }
      __syncwarp();
    }
#pragma unroll
    for (int t = 0; t < T; t++) {
      if (lane + 32 * t < NS) {
        if (rowslot) {                                            // packed rows: row i of block k at the slot of the pair (k, i)
          constexpr int NPP = NP + (NP & 1);
          const int si = __ldg(rowslot + k * NP + ei[t]), sj = __ldg(rowslot + k * NP + ej[t]);
          if (si >= 0) mats[(int64_t)si * NPP + ej[t]] = acc[t];
          if (sj >= 0) mats[(int64_t)sj * NPP + ei[t]] = acc[t];
        } else {
          mats[k * NE + ei[t] * NP + ej[t]] = acc[t];
          mats[k * NE + ej[t] * NP + ei[t]] = acc[t];
        }
      }
    }
  }
}

template <int DIM, int ORDER>
static int launch_fpc(const ngsx_mesh &M, const DeviceData *dd, const double4 *rule, int nr, double coef, int hpow, int64_t np,
                      const int32_t *patch_els, const uint8_t *cidx, int ndp, double *mats, const int32_t *rowslot, cudaStream_t st) {
  constexpr int ND = H1<DIM, ORDER>::ND, NFD = DIM == 2 ? ORDER + 1 : (ORDER == 1 ? 3 : (ORDER == 2 ? 6 : 10)), NP = 2 * ND - NFD;
  if (ndp != NP) { set_error("facetpatch_cmat: ndp does not match the merged dof count of this element type"); return NGSX_ERR_ARG; }
  size_t smem = (size_t)nr * 32 + (size_t)FP_WARPS * (32 * NP + 32) * 8;
  auto kern = k_facetpatch_c<DIM, ORDER>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t need = (np + FP_WARPS - 1) / FP_WARPS;
  unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)dd->sm_count * 8));
  kern<<<grid, FP_WARPS * 32, smem, st>>>(M, rule, nr, coef, hpow, np, patch_els, cidx, mats, rowslot);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

// ---- merged patch blocks on a DEFORMED mesh (row f-4) ---------------------------------------------------------------------
// Both elements of a patch are isoparametric (x = sum_i c_i phi_i, order DORD); a rule point of one element is pulled into the
// other by Newton's method exactly as the reference's MapPatchIntegrationPoint does (xfem/symboliccutbfi.cpp:1113-1199): start
// from the affine map through the vertices, at most 200 steps, stop when the residual before the last update was
// <= 1e-12 sqrt(det F_from) (NaN, i.e. one step, on a negatively oriented element), else the "low order guess".  The weight
// is mir1.GetWeight(): for a point that came from element 2 the reference's measure_2 w / measure_1(previous iterate) times
// measure_1(final iterate); the coefficient h = specialcf.mesh_size is evaluated on mir1 per point (:1359, 1365).
struct DefoArg { const int32_t *el2dof; const double *coef; };

template <int DIM, int DORD>
__device__ __forceinline__ void load_gc(const ngsx_mesh &M, const DefoArg &D, int64_t e, int lane, double *gc) {
  constexpr int NDG = H1<DIM, DORD>::ND;
  for (int t = lane; t < NDG * DIM; t += 32) {
    const int i = t / DIM, d = t % DIM;
    double c = __ldg(D.coef + (int64_t)__ldg(D.el2dof + e * NDG + i) * DIM + d);
    if (i <= DIM) c += __ldg(M.vert + (int64_t)__ldg(M.el2v + e * (DIM + 1) + i) * DIM + d);
    gc[t] = c;
  }
}

template <int DIM> __device__ __forceinline__ void lam_of(const double *xi, double *lam) {
  double last = 1.0;
  lam[3] = 0.0;
#pragma unroll
  for (int k = 0; k < DIM; k++) { lam[k] = xi[k]; last -= xi[k]; }
  lam[DIM] = last;
}

template <int DIM, int ORDER, int DORD>
__global__ void __launch_bounds__(FP_WARPS * 32) k_facetpatch_cd(ngsx_mesh M, DefoArg D, const double4 *__restrict__ rule, int nr, double coef,
                                                                int hpow, int64_t npatch, const int32_t *__restrict__ patch_els,
                                                                const uint8_t *__restrict__ cidx, double *__restrict__ mats,
                                                                const int32_t *__restrict__ rowslot) {
  constexpr int ND = H1<DIM, ORDER>::ND, NFD = DIM == 2 ? (ORDER == 1 ? 2 : 3) : (ORDER == 1 ? 3 : 6), NDG = H1<DIM, DORD>::ND;
  constexpr int NP = 2 * ND - NFD, NE = NP * NP, NS = NP * (NP + 1) / 2, T = (NS + 31) / 32;
  constexpr int PER_WARP = 32 * NP + 32 + 2 * NDG * DIM;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double4 *srule = reinterpret_cast<double4 *>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *stage = reinterpret_cast<double *>(smem_raw + (size_t)nr * 32) + warp * PER_WARP;
  double *facq = stage + 32 * NP;
  double *gc0 = facq + 32, *gc1 = gc0 + NDG * DIM;
  for (int i = threadIdx.x; i < nr; i += blockDim.x) srule[i] = rule[i];
  __syncthreads();
  int ei[T], ej[T];
#pragma unroll
  for (int t = 0; t < T; t++) {
    int idx = lane + 32 * t, i = 0;
    if (idx >= NS) idx = 0;
    while (idx >= NP - i) { idx -= NP - i; i++; }
    ei[t] = i; ej[t] = i + idx;
  }
  const int64_t nwarps = (int64_t)gridDim.x * FP_WARPS;
  for (int64_t k = (int64_t)blockIdx.x * FP_WARPS + warp; k < npatch; k += nwarps) {
    __syncwarp();
    load_gc<DIM, DORD>(M, D, patch_els[2 * k], lane, gc0);
    load_gc<DIM, DORD>(M, D, patch_els[2 * k + 1], lane, gc1);
    __syncwarp();
    const uint8_t *ci = cidx + k * 2 * ND;
    double acc[T];
#pragma unroll
    for (int t = 0; t < T; t++) acc[t] = 0.0;
    const int ntot = 2 * nr;
    for (int q0 = 0; q0 < ntot; q0 += 32) {
      const int nq = min(32, ntot - q0);
      if (lane < nq) {
        const int q = q0 + lane, side = q >= nr ? 1 : 0;
        const double4 r = srule[q - side * nr];
        const double *gf = side ? gc1 : gc0, *gt = side ? gc0 : gc1;
        double xif[3] = {r.x, r.y, r.z}, lamf[4];
        lam_of<DIM>(xif, lamf);
        double x[DIM], gl[DIM + 1][DIM], detf;
        iso_point<DIM, DORD>(GcWarp{gf}, lamf, x, gl, detf);
        const double tol = 1e-12 * sqrt(detf);                          // NaN on a negatively oriented element: one Newton step
        const double wfrom = fabs(detf) * r.w;
        // start: affine map through the vertices of the target element (its vertex coefficients: edge functions vanish there)
        double A[DIM][DIM], Ai[DIM][DIM], detA;
#pragma unroll
        for (int rr = 0; rr < DIM; rr++)
#pragma unroll
          for (int d = 0; d < DIM; d++) A[rr][d] = gt[d * DIM + rr] - gt[DIM * DIM + rr];
        invert<DIM>(A, Ai, detA);
        double xi0[3] = {0.0, 0.0, 0.0}, xit[3] = {0.0, 0.0, 0.0};
#pragma unroll
        for (int kk = 0; kk < DIM; kk++) {
          double v = 0.0;
#pragma unroll
          for (int rr = 0; rr < DIM; rr++) v += Ai[kk][rr] * (x[rr] - gt[DIM * DIM + rr]);
          xi0[kk] = v; xit[kk] = v;
        }
        int its = 0;
        double wlast = 0.0, nrm = 0.0;
        while (its == 0 || (nrm > tol && its < 200)) {
          double lamt[4], xq[DIM], glq[DIM + 1][DIM], dq;
          lam_of<DIM>(xit, lamt);
          iso_point<DIM, DORD>(GcWarp{gt}, lamt, xq, glq, dq);
          double diff[DIM];
          nrm = 0.0;
#pragma unroll
          for (int rr = 0; rr < DIM; rr++) { diff[rr] = x[rr] - xq[rr]; nrm += diff[rr] * diff[rr]; }
          nrm = sqrt(nrm);
#pragma unroll
          for (int kk = 0; kk < DIM; kk++) {
            double v = 0.0;
#pragma unroll
            for (int rr = 0; rr < DIM; rr++) v += glq[kk][rr] * diff[rr];
            xit[kk] += v;
          }
          its++;
          wlast = fabs(dq);
        }
        double wto = wfrom / wlast;
        if (its >= 200) {                                               // "taking a low order guess"
#pragma unroll
          for (int kk = 0; kk < 3; kk++) xit[kk] = xi0[kk];
          wto = wfrom / fabs(detA);
        }
        // mir1 = element 0 at its point of the pair; weight mir1.GetWeight()
        double lam0[4], lam1[4], det0 = detf, w1 = r.w;
        if (side == 0) { lam_of<DIM>(xif, lam0); lam_of<DIM>(xit, lam1); }
        else {
          lam_of<DIM>(xit, lam0); lam_of<DIM>(xif, lam1);
          double xq[DIM], glq[DIM + 1][DIM];
          iso_point<DIM, DORD>(GcWarp{gc0}, lam0, xq, glq, det0);
          w1 = wto;
        }
        const double m0 = fabs(det0);
        const double h = DIM == 2 ? sqrt(m0) : cbrt(m0);
        double *jc = stage + lane * NP;
        for (int c = 0; c < NP; c++) jc[c] = 0.0;
        for (int i = 0; i < ND; i++) {
          jc[ci[i]] += h1_shape<DIM, ORDER>(i, lam0);
          jc[ci[ND + i]] -= h1_shape<DIM, ORDER>(i, lam1);
        }
        facq[lane] = coef * ipow(h, hpow) * (m0 * w1);
      }
      __syncwarp();
      for (int q = 0; q < nq; q++) {
        const double f = facq[q];
#pragma unroll
        for (int t = 0; t < T; t++) acc[t] += f * stage[q * NP + ei[t]] * stage[q * NP + ej[t]];
      }
      __syncwarp();
    }
#pragma unroll
    for (int t = 0; t < T; t++) {
      if (lane + 32 * t < NS) {
        if (rowslot) {
          constexpr int NPP = NP + (NP & 1);
          const int si = __ldg(rowslot + k * NP + ei[t]), sj = __ldg(rowslot + k * NP + ej[t]);
          if (si >= 0) mats[(int64_t)si * NPP + ej[t]] = acc[t];
          if (sj >= 0) mats[(int64_t)sj * NPP + ei[t]] = acc[t];
        } else {
          mats[k * NE + ei[t] * NP + ej[t]] = acc[t];
          mats[k * NE + ej[t] * NP + ei[t]] = acc[t];
        }
      }
    }
  }
}

template <int DIM, int ORDER, int DORD>
static int launch_fpcd(const ngsx_mesh &M, const DefoArg &D, const DeviceData *dd, const double4 *rule, int nr, double coef, int hpow, int64_t np,
                       const int32_t *patch_els, const uint8_t *cidx, int ndp, double *mats, const int32_t *rowslot, cudaStream_t st) {
  constexpr int ND = H1<DIM, ORDER>::ND, NFD = DIM == 2 ? (ORDER == 1 ? 2 : 3) : (ORDER == 1 ? 3 : 6), NP = 2 * ND - NFD;
  constexpr int NDG = H1<DIM, DORD>::ND;
  if (ndp != NP) { set_error("facetpatch_cmat_deformed: ndp does not match the merged dof count of this element type"); return NGSX_ERR_ARG; }
  size_t smem = (size_t)nr * 32 + (size_t)FP_WARPS * (32 * NP + 32 + 2 * NDG * DIM) * 8;
  auto kern = k_facetpatch_cd<DIM, ORDER, DORD>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t need = (np + FP_WARPS - 1) / FP_WARPS;
  unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)dd->sm_count * 8));
  kern<<<grid, FP_WARPS * 32, smem, st>>>(M, D, rule, nr, coef, hpow, np, patch_els, cidx, mats, rowslot);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

}  // namespace ngsx

using namespace ngsx;

extern "C" int ngsx_facetpatch_elmat_h1(const ngsx_mesh *mesh, int fe_order, double coef, int hpow, int64_t npatch,
                                        const int32_t *patch_els, double *elmats, void *stream) {
  NGSX_REQUIRE(mesh && (mesh->dim == 2 || mesh->dim == 3) && mesh->vert && mesh->el2v, "facetpatch: bad mesh");
  NGSX_REQUIRE(npatch >= 0 && (npatch == 0 || (patch_els && elmats)), "facetpatch: bad arguments");
  if (fe_order < 1 || fe_order > 2) { set_error("facetpatch: H1 order outside the CUDA path (1, 2)"); return NGSX_ERR_UNSUPPORTED; }
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  const double4 *rule; int nr;
  if (!rule_lookup(dd, mesh->dim, 2 * fe_order, &rule, &nr)) { set_error("facetpatch: rule not tabulated"); return NGSX_ERR_UNSUPPORTED; }
  if (npatch == 0) return NGSX_OK;
  cudaStream_t st = as_stream(stream);
  switch (mesh->dim * 10 + fe_order) {
    case 21: return launch_fp<2, 1>(*mesh, dd, rule, nr, coef, hpow, npatch, patch_els, elmats, st);
    case 22: return launch_fp<2, 2>(*mesh, dd, rule, nr, coef, hpow, npatch, patch_els, elmats, st);
    case 31: return launch_fp<3, 1>(*mesh, dd, rule, nr, coef, hpow, npatch, patch_els, elmats, st);
    case 32: return launch_fp<3, 2>(*mesh, dd, rule, nr, coef, hpow, npatch, patch_els, elmats, st);
  }
  return NGSX_ERR_UNSUPPORTED;
}

extern "C" int ngsx_facetpatch_table_h1(int64_t npatch, const int32_t *patch_els, int32_t nd, const int32_t *el2dof, int32_t ndp,
                                        int32_t *table, uint8_t *cidx, void *stream) {
  NGSX_REQUIRE(npatch >= 0 && nd > 0 && nd <= 64 && ndp >= nd && ndp <= 2 * nd, "facetpatch_table: bad arguments");
  if (npatch == 0) return NGSX_OK;
  NGSX_REQUIRE(patch_els && el2dof && table && cidx, "facetpatch_table: NULL array");
  cudaStream_t st = as_stream(stream);
  int *bad = nullptr;
  NGSX_CUDA(cudaMallocAsync(&bad, sizeof(int), st));
  NGSX_CUDA(cudaMemsetAsync(bad, 0, sizeof(int), st));
  k_fp_table<<<(unsigned)((npatch + 127) / 128), 128, 0, st>>>(npatch, patch_els, nd, el2dof, ndp, table, cidx, bad);
  g_launches.fetch_add(1);
  int h_bad = 0;
  cudaError_t e = cudaMemcpyAsync(&h_bad, bad, sizeof(int), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFreeAsync(bad, st);
  NGSX_CUDA(e);
  if (h_bad) { set_error("facetpatch_table: a patch does not share exactly 2 nd - ndp dofs (non-conforming space or wrong ndp)"); return NGSX_ERR_ARG; }
  return NGSX_OK;
}

extern "C" int ngsx_facetpatch_cmat_h1(const ngsx_mesh *mesh, int fe_order, double coef, int hpow, int64_t npatch,
                                       const int32_t *patch_els, const uint8_t *cidx, int32_t ndp, double *mats, void *stream) {
  NGSX_REQUIRE(mesh && (mesh->dim == 2 || mesh->dim == 3) && mesh->vert && mesh->el2v, "facetpatch_cmat: bad mesh");
  NGSX_REQUIRE(npatch >= 0 && (npatch == 0 || (patch_els && cidx && mats)), "facetpatch_cmat: bad arguments");
  if (fe_order < 1 || fe_order > 3) { set_error("facetpatch_cmat: H1 order outside the CUDA path (1, 2, 3)"); return NGSX_ERR_UNSUPPORTED; }
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  const double4 *rule; int nr;
  if (!rule_lookup(dd, mesh->dim, 2 * fe_order, &rule, &nr)) { set_error("facetpatch_cmat: rule not tabulated"); return NGSX_ERR_UNSUPPORTED; }
  const int32_t *rowslot = take_packed_output();
  if (npatch == 0) return NGSX_OK;
  cudaStream_t st = as_stream(stream);
  switch (mesh->dim * 10 + fe_order) {
    case 21: return launch_fpc<2, 1>(*mesh, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 22: return launch_fpc<2, 2>(*mesh, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 31: return launch_fpc<3, 1>(*mesh, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 32: return launch_fpc<3, 2>(*mesh, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 23: return launch_fpc<2, 3>(*mesh, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 33: return launch_fpc<3, 3>(*mesh, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);   // 30 merged dofs: dense blocks only (the gather's descriptors end at 26)
  }
  return NGSX_ERR_UNSUPPORTED;
}

extern "C" int ngsx_facetpatch_cmat_h1_deformed(const ngsx_mesh *mesh, const ngsx_deformation *defo, int fe_order, double coef, int hpow,
                                                int64_t npatch, const int32_t *patch_els, const uint8_t *cidx, int32_t ndp, double *mats,
                                                void *stream) {
  NGSX_REQUIRE(mesh && (mesh->dim == 2 || mesh->dim == 3) && mesh->vert && mesh->el2v, "facetpatch_cmat_deformed: bad mesh");
  NGSX_REQUIRE(defo && defo->el2dof && defo->coef, "facetpatch_cmat_deformed: bad deformation");
  NGSX_REQUIRE(npatch >= 0 && (npatch == 0 || (patch_els && cidx && mats)), "facetpatch_cmat_deformed: bad arguments");
  if (fe_order < 1 || fe_order > 2 || defo->order < 1 || defo->order > 2) {
    set_error("facetpatch_cmat_deformed: H1 / deformation order outside the CUDA path (1, 2)");
    return NGSX_ERR_UNSUPPORTED;
  }
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  const double4 *rule; int nr;
  if (!rule_lookup(dd, mesh->dim, 2 * fe_order, &rule, &nr)) { set_error("facetpatch_cmat_deformed: rule not tabulated"); return NGSX_ERR_UNSUPPORTED; }
  const int32_t *rowslot = take_packed_output();
  if (npatch == 0) return NGSX_OK;
  cudaStream_t st = as_stream(stream);
  const DefoArg D{defo->el2dof, defo->coef};
  switch (mesh->dim * 100 + fe_order * 10 + defo->order) {
    case 211: return launch_fpcd<2, 1, 1>(*mesh, D, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 212: return launch_fpcd<2, 1, 2>(*mesh, D, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 221: return launch_fpcd<2, 2, 1>(*mesh, D, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 222: return launch_fpcd<2, 2, 2>(*mesh, D, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 311: return launch_fpcd<3, 1, 1>(*mesh, D, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 312: return launch_fpcd<3, 1, 2>(*mesh, D, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 321: return launch_fpcd<3, 2, 1>(*mesh, D, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
    case 322: return launch_fpcd<3, 2, 2>(*mesh, D, dd, rule, nr, coef, hpow, npatch, patch_els, cidx, ndp, mats, rowslot, st);
  }
  return NGSX_ERR_UNSUPPORTED;
}
