// ngsx_spacetime.cu -- (f-3) space-time cut quadrature: SpaceTimeCutIntegrationRule (cutint/spacetimecutrule.cpp:98-196) and
// IntegrateX with time_order >= 0 (cutint/python_cutint.cpp:20-72 -> CreateCutIntegrationRule, cutint/xintegration.cpp:38-58).
//
// The level set is a space-time function: P1 in space, nodal of order 0..2 in the reference time (NodalTimeFE nodes 0 | 0,1 |
// 0,1/2,1, spacetime/SpaceTimeFE.cpp:150-160), stored time-major lset_st[k * nv + vertex] (SpaceTimeFESpace::GetDofNrs,
// spacetime/SpaceTimeFESpace.cpp:109-122).  Per element: the zeros in time of every vertex trace (root_finding, closed forms of
// cutint/spacetimecutrule.cpp:6-45) split [0,1] into sub-intervals; on each, every Gauss point in time freezes a P1 level set
// (|value| < 1e-14 -> 1e-14) whose straight-cut rule -- or the static rule on uncut elements of the domain -- is emitted with
// weight w_space * w_time * (t1 - t0) and the reference time of the point.
// One thread per element walks this nest (st_walk) with a functor per emitted point: counting, emission at scan offsets in
// the reference's order, or integration of a polynomial in (x, tref) with a fixed-shape reduction (bitwise reproducible).
#include <algorithm>
#include "ngsx_host.cuh"

namespace ngsx {

constexpr int ST_MAXCP = 12;      // 2 + 2 roots per vertex of a tet

// zeros in (0,1) of one vertex trace; same expressions as the reference (including a == 0 on an order-2 element: inf / nan)
__device__ __forceinline__ int st_roots(int ntd, const double *li, double *roots) {
  int n = 0;
  if (ntd == 2) {
    if ((li[0] >= 0) != (li[1] >= 0)) roots[n++] = -li[0] / (li[1] - li[0]);
  } else if (ntd == 3) {
    const double c = li[0], a = 2 * li[0] + 2 * li[2] - 4. * li[1], b = li[2] - a - c;
    const double w = b * b - 4 * a * c;
    if (fabs(w) < 1e-12) {
      const double x = -b / (2 * a);
      if (x < 1 && x > 0) roots[n++] = x;
    } else {
      const double x1 = (b + sqrt(w)) / (-2.0 * a), x2 = (b - sqrt(w)) / (-2.0 * a);
      if (x1 < 1 && x1 > 0) roots[n++] = x1;
      if (x2 < 1 && x2 > 0) roots[n++] = x2;
    }
  }
  return n;
}

__device__ __forceinline__ void st_time_shape(int ntd, double t, double *sh) {
  if (ntd == 1) { sh[0] = 1.0; sh[1] = sh[2] = 0.0; }
  else if (ntd == 2) { sh[0] = 1.0 - t; sh[1] = t; sh[2] = 0.0; }
  else { sh[0] = (1 - t) * (1 - 2 * t); sh[1] = 4 * t * (1 - t); sh[2] = t * (2 * t - 1); }
}

struct StArgs {
  ngsx_mesh M;
  const double *lset_st;
  int ntd, dt;
  const CutTable *tab;
  const double4 *rule_t; int nrt;
  const double4 *rule_vol; int nr_vol;
  const double4 *rule_if; int nr_if;
};

template <int DIM>
__device__ __forceinline__ double st_interface_scale(const ElGeom<DIM> &G, const double *ls) {
  double g[DIM], n2 = 0.0;
#pragma unroll
  for (int k = 0; k < DIM; k++) { g[k] = ls[k] - ls[DIM]; n2 += g[k] * g[k]; }
  const double inv = 1.0 / sqrt(n2);
  double s2 = 0.0;
#pragma unroll
  for (int r = 0; r < DIM; r++) {
    double v = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; k++) v += G.Jinv[k][r] * (g[k] * inv);
    s2 += v * v;
  }
  return sqrt(s2);
}

// emit(xi[3], t, w): every point of the element's space-time rule in the reference's order.  POINTS = false: emit is called
// once per spatial rule with the number of points only (counting).
template <int DIM, bool POINTS, class F>
__device__ __forceinline__ void st_walk(const StArgs &A, int64_t e, F &emit) {
  constexpr int NV = DIM + 1;
  int v[NV];
  double lst[3][NV], p[NV][DIM];
#pragma unroll
  for (int k = 0; k < NV; k++) {
    v[k] = __ldg(A.M.el2v + e * NV + k);
    if (POINTS) {
#pragma unroll
      for (int d = 0; d < DIM; d++) p[k][d] = __ldg(A.M.vert + (int64_t)v[k] * DIM + d);
    }
  }
  for (int k = 0; k < 3; k++)
#pragma unroll
    for (int i = 0; i < NV; i++) lst[k][i] = k < A.ntd ? __ldg(A.lset_st + (int64_t)k * A.M.nv + v[i]) : 0.0;
  double cp[ST_MAXCP];
  int ncp = 0;
  cp[ncp++] = 0.0; cp[ncp++] = 1.0;
#pragma unroll
  for (int i = 0; i < NV; i++) {
    double li[3] = {lst[0][i], lst[1][i], lst[2][i]}, r[2];
    const int nr = st_roots(A.ntd, li, r);
    for (int k = 0; k < nr; k++) cp[ncp++] = r[k];
  }
  for (int i = 1; i < ncp; i++) {                    // insertion sort (std::sort in the reference: values only)
    const double val = cp[i];
    int j = i - 1;
    while (j >= 0 && cp[j] > val) { cp[j + 1] = cp[j]; j--; }
    cp[j + 1] = val;
  }
  ElGeom<DIM> G;
  if (POINTS && A.dt == IF) element_geometry<DIM>(p, G);
  for (int i = 0; i + 1 < ncp; i++) {
    const double t0 = cp[i], t1 = cp[i + 1];
    for (int q = 0; q < A.nrt; q++) {
      const double4 rt = A.rule_t[q];
      const double t = t0 + rt.x * (t1 - t0);
      double sh[3], ls[4];
      st_time_shape(A.ntd, t, sh);
#pragma unroll
      for (int k = 0; k < NV; k++) {
        double d = lst[0][k] * sh[0];
        if (A.ntd > 1) d += lst[1][k] * sh[1];
        if (A.ntd > 2) d += lst[2][k] * sh[2];
        if (fabs(d) < 1e-14) d = 1e-14;
        ls[k] = d;
      }
      int mask;
      const int dom = sign_test<DIM>(ls, mask);
      const double wt = rt.w * (t1 - t0);
      if (dom == IF) {
        const CutEntry E = A.tab->e[DIM - 2][A.dt][mask];
        const double4 *rule = A.dt == IF ? A.rule_if : A.rule_vol;
        const int nr = A.dt == IF ? A.nr_if : A.nr_vol;
        if constexpr (!POINTS) {
          emit(nullptr, t, (double)(E.ns * nr));
        } else {
          const double scale = A.dt == IF ? st_interface_scale<DIM>(G, ls) : 1.0;
          for (int s = 0; s < E.ns; s++) {
            double P[4][3];
            for (int k = 0; k < E.npt; k++) coded_point<DIM>(E.v[s][k], ls, P[k]);
            const double fac = simplex_factor(E.npt, P) * scale;
            for (int qq = 0; qq < nr; qq++) {
              const double4 r = rule[qq];
              double xi[3];
              simplex_point(E.npt, P, r, xi);
              emit(xi, t, (r.w * fac) * wt);
            }
          }
        }
      } else if (dom == A.dt) {
        if constexpr (!POINTS) {
          emit(nullptr, t, (double)A.nr_vol);
        } else {
          for (int qq = 0; qq < A.nr_vol; qq++) {
            const double4 r = A.rule_vol[qq];
            const double xi[3] = {r.x, r.y, r.z};
            emit(xi, t, r.w * wt);
          }
        }
      }
    }
  }
}

struct CountF {
  int n = 0;
  __device__ __forceinline__ void operator()(const double *, double, double cnt) { n += (int)cnt; }
};

template <int DIM>
__global__ void __launch_bounds__(128) k_st_count(StArgs A, int32_t *__restrict__ npts) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= A.M.ne) return;
  CountF f;
  st_walk<DIM, false>(A, e, f);
  npts[e] = f.n;
}

struct FillF {
  double *pts3, *tq, *wts;
  int64_t o;
  __device__ __forceinline__ void operator()(const double *xi, double t, double w) {
    pts3[3 * o] = xi[0]; pts3[3 * o + 1] = xi[1]; pts3[3 * o + 2] = xi[2];
    tq[o] = t; wts[o] = w;
    o++;
  }
};

template <int DIM>
__global__ void __launch_bounds__(128) k_st_fill(StArgs A, const int64_t *__restrict__ offs, double *__restrict__ pts3, double *__restrict__ tq,
                                                double *__restrict__ wts) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= A.M.ne) return;
  FillF f{pts3, tq, wts, offs[e]};
  st_walk<DIM, true>(A, e, f);
}

struct StPoly { int n; double c[NGSX_POLY_MAXTERMS]; int ex[NGSX_POLY_MAXTERMS][4]; };

__device__ __forceinline__ double st_ipow(double x, int p) {
  double r = 1.0;
  for (int k = 0; k < p; k++) r *= x;
  return r;
}

template <int DIM>
struct IntF {
  const StPoly *f;
  const double (*p)[DIM];
  double sum = 0.0;
  __device__ __forceinline__ void operator()(const double *xi, double t, double w) {
    double lam_last = 1.0, x[3] = {0.0, 0.0, 0.0};
#pragma unroll
    for (int k = 0; k < DIM; k++) lam_last -= xi[k];
#pragma unroll
    for (int d = 0; d < DIM; d++) {
      double val = lam_last * p[DIM][d];
#pragma unroll
      for (int k = 0; k < DIM; k++) val += xi[k] * p[k][d];
      x[d] = val;
    }
    double fx = 0.0;
    for (int m = 0; m < f->n; m++)
      fx += f->c[m] * st_ipow(x[0], f->ex[m][0]) * st_ipow(x[1], f->ex[m][1]) * st_ipow(x[2], f->ex[m][2]) * st_ipow(t, f->ex[m][3]);
    sum += w * fx;
  }
};

template <int DIM>
__global__ void __launch_bounds__(128) k_st_integrate(StArgs A, StPoly f, double *__restrict__ partial) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double val = 0.0;
  if (e < A.M.ne) {
    double p[DIM + 1][DIM];
#pragma unroll
    for (int k = 0; k <= DIM; k++) {
      const int v = __ldg(A.M.el2v + e * (DIM + 1) + k);
#pragma unroll
      for (int d = 0; d < DIM; d++) p[k][d] = __ldg(A.M.vert + (int64_t)v * DIM + d);
    }
    ElGeom<DIM> G;
    element_geometry<DIM>(p, G);
    IntF<DIM> F{&f, p};
    st_walk<DIM, true>(A, e, F);
    val = fabs(G.det) * F.sum;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
  if ((threadIdx.x & 31) == 0) partial[e >> 5] = val;
}

// fixed-shape reduction of the warp partials (same scheme as ngsx_integrate)
__global__ void __launch_bounds__(256) k_st_reduce(int64_t n, const double *__restrict__ in, double *__restrict__ partial) {
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

__global__ void __launch_bounds__(1024) k_st_final(int64_t n, const double *__restrict__ partial, double *__restrict__ out) {
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

static int st_args(const ngsx_mesh *mesh, const double *lset_st, int ntd, int dt, int order_space, int order_time, StArgs &A) {
  NGSX_REQUIRE(mesh && (mesh->dim == 2 || mesh->dim == 3) && mesh->vert && mesh->el2v && lset_st, "space-time rule: bad mesh / level set");
  NGSX_REQUIRE(dt >= 0 && dt <= 2, "space-time rule: domain_type must be NEG, POS or IF");
  if (ntd < 1 || ntd > 3) { set_error("space-time rule: time element of order > 2 (bisection root finding) is outside the CUDA path"); return NGSX_ERR_UNSUPPORTED; }
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  A.M = *mesh; A.lset_st = lset_st; A.ntd = ntd; A.dt = dt; A.tab = dd->cut_table;
  if (!rule_lookup(dd, 1, order_time, &A.rule_t, &A.nrt) || !rule_lookup(dd, mesh->dim, order_space, &A.rule_vol, &A.nr_vol) ||
      !rule_lookup(dd, mesh->dim - 1, order_space, &A.rule_if, &A.nr_if)) {
    set_error("space-time rule: integration order beyond the tabulated reference rules");
    return NGSX_ERR_UNSUPPORTED;
  }
  return NGSX_OK;
}

}  // namespace ngsx

using namespace ngsx;

extern "C" {

int ngsx_st_rule_count(const ngsx_mesh *mesh, const double *lset_st, int ntime_dofs, int dt, int order_space, int order_time,
                       int32_t *npts, void *stream) {
  StArgs A{};
  if (int rc = st_args(mesh, lset_st, ntime_dofs, dt, order_space, order_time, A)) return rc;
  if (mesh->ne == 0) return NGSX_OK;
  NGSX_REQUIRE(npts, "st_rule_count: npts is NULL");
  const unsigned nb = (unsigned)((mesh->ne + 127) / 128);
  if (mesh->dim == 2) k_st_count<2><<<nb, 128, 0, as_stream(stream)>>>(A, npts);
  else k_st_count<3><<<nb, 128, 0, as_stream(stream)>>>(A, npts);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_st_rule_fill(const ngsx_mesh *mesh, const double *lset_st, int ntime_dofs, int dt, int order_space, int order_time,
                      const int64_t *offs, double *pts3, double *tref, double *wts, void *stream) {
  StArgs A{};
  if (int rc = st_args(mesh, lset_st, ntime_dofs, dt, order_space, order_time, A)) return rc;
  if (mesh->ne == 0) return NGSX_OK;
  NGSX_REQUIRE(offs && pts3 && tref && wts, "st_rule_fill: NULL array");
  const unsigned nb = (unsigned)((mesh->ne + 127) / 128);
  if (mesh->dim == 2) k_st_fill<2><<<nb, 128, 0, as_stream(stream)>>>(A, offs, pts3, tref, wts);
  else k_st_fill<3><<<nb, 128, 0, as_stream(stream)>>>(A, offs, pts3, tref, wts);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_st_integrate(const ngsx_mesh *mesh, const double *lset_st, int ntime_dofs, int dt, int order_space, int order_time,
                      const ngsx_poly *h_f, const int32_t *h_texp, double *out, void *stream) {
  StArgs A{};
  if (int rc = st_args(mesh, lset_st, ntime_dofs, dt, order_space, order_time, A)) return rc;
  NGSX_REQUIRE(h_f && out && h_f->nterms >= 0 && h_f->nterms <= NGSX_POLY_MAXTERMS, "st_integrate: bad integrand");
  cudaStream_t st = as_stream(stream);
  if (mesh->ne == 0) { NGSX_CUDA(cudaMemsetAsync(out, 0, sizeof(double), st)); return NGSX_OK; }
  StPoly f;
  f.n = h_f->nterms;
  for (int m = 0; m < f.n; m++) {
    f.c[m] = h_f->coef[m];
    for (int d = 0; d < 3; d++) f.ex[m][d] = h_f->ex[m][d];
    f.ex[m][3] = h_texp ? h_texp[m] : 0;
    NGSX_REQUIRE(f.ex[m][3] >= 0, "st_integrate: negative power of tref");
  }
  const unsigned nb = (unsigned)((mesh->ne + 127) / 128);
  const int64_t nred = (int64_t)nb * 4;
  const unsigned nmid = (unsigned)std::min<int64_t>((nred + 255) / 256, 296);
  double *partial = nullptr;
  NGSX_CUDA(cudaMallocAsync(&partial, (nred + nmid) * sizeof(double), st));
  if (mesh->dim == 2) k_st_integrate<2><<<nb, 128, 0, st>>>(A, f, partial);
  else k_st_integrate<3><<<nb, 128, 0, st>>>(A, f, partial);
  cudaError_t e1 = cudaGetLastError();
  k_st_reduce<<<nmid, 256, 0, st>>>(nred, partial, partial + nred);
  k_st_final<<<1, 1024, 0, st>>>(nmid, partial + nred, out);
  g_launches.fetch_add(3);
  cudaError_t e2 = cudaGetLastError();
  cudaFreeAsync(partial, st);
  NGSX_CUDA(e1);
  NGSX_CUDA(e2);
  return NGSX_OK;
}

}  // extern "C"
