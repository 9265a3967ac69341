// ngsx_spacetime.cu -- (f-3) space-time cut quadrature: SpaceTimeCutIntegrationRule (cutint/spacetimecutrule.cpp:98-196) and
// IntegrateX with time_order >= 0 (cutint/python_cutint.cpp:20-72 -> CreateCutIntegrationRule, cutint/xintegration.cpp:38-58).
//
// The level set is a space-time function: P1 in space, nodal of order 0..5 in the reference time (NodalTimeFE: Gauss-Lobatto nodes,
// spacetime/SpaceTimeFE.cpp:150-173; closed-form roots for orders 1, 2, the reference's subdivision + secant / bisection search
// above, cutint/spacetimecutrule.cpp:46-95), stored time-major lset_st[k * nv + vertex] (SpaceTimeFESpace::GetDofNrs,
// spacetime/SpaceTimeFESpace.cpp:109-122).  Per element: the zeros in time of every vertex trace (root_finding, closed forms of
// cutint/spacetimecutrule.cpp:6-45) split [0,1] into sub-intervals; on each, every Gauss point in time freezes a P1 level set
// (|value| < 1e-14 -> 1e-14) whose straight-cut rule -- or the static rule on uncut elements of the domain -- is emitted with
// weight w_space * w_time * (t1 - t0) and the reference time of the point.
// One thread per element walks this nest (st_walk) with a functor per emitted point: counting, emission at scan offsets in
// the reference's order, or integration of a polynomial in (x, tref) with a fixed-shape reduction (bitwise reproducible).
#include <algorithm>
#include "ngsx_host.cuh"

namespace ngsx {

constexpr int ST_MAXTD = 6;                          // time dofs of the nodal element (order <= 5, as NodalTimeFE::CalcInterpolationPoints)
constexpr int ST_MAXCP = 2 + 4 * (ST_MAXTD - 1);     // 0, 1 and at most `order` roots per vertex of a tet

// NodalTimeFE::CalcInterpolationPoints (spacetime/SpaceTimeFE.cpp:150-173), same expressions
__device__ __forceinline__ void st_nodes(int ntd, double *nd) {
  nd[0] = 0.0;
  if (ntd == 2) nd[1] = 1.0;
  else if (ntd == 3) { nd[1] = 0.5; nd[2] = 1.0; }
  else if (ntd == 4) { nd[1] = 0.5 * (1.0 - 1.0 / sqrt(5.0)); nd[2] = 0.5 * (1.0 + 1.0 / sqrt(5.0)); nd[3] = 1.0; }
  else if (ntd == 5) { nd[1] = 0.5 * (1.0 - sqrt(3.0 / 7.0)); nd[2] = 0.5; nd[3] = 0.5 * (1.0 + sqrt(3.0 / 7.0)); nd[4] = 1.0; }
  else if (ntd == 6) {
    nd[1] = 0.5 * (1.0 - sqrt(1.0 / 3.0 + 2.0 * sqrt(7.0) / 21.0));
    nd[2] = 0.5 * (1.0 - sqrt(1.0 / 3.0 - 2.0 * sqrt(7.0) / 21.0));
    nd[3] = 0.5 * (1.0 + sqrt(1.0 / 3.0 - 2.0 * sqrt(7.0) / 21.0));
    nd[4] = 0.5 * (1.0 + sqrt(1.0 / 3.0 + 2.0 * sqrt(7.0) / 21.0));
    nd[5] = 1.0;
  }
}

// NodalTimeFE::CalcShape: Lagrange_Pol (spacetime/SpaceTimeFE.hpp:86-95), the product in the reference's order
__device__ __forceinline__ void st_lagrange(int ntd, const double *nd, double x, double *sh) {
  for (int i = 0; i < ntd; i++) {
    double r = 1.0;
    for (int j = 0; j < ntd; j++)
      if (j != i) r *= (x - nd[j]) / (nd[i] - nd[j]);
    sh[i] = r;
  }
}

// d/dx of the Lagrange polynomials (NodalTimeFE::CalcDShape: AutoDiff through the same product)
__device__ __forceinline__ void st_dlagrange(int ntd, const double *nd, double x, double *dsh) {
  for (int i = 0; i < ntd; i++) {
    double val = 1.0, der = 0.0;                      // product rule, factor by factor like AutoDiff<1>
    for (int j = 0; j < ntd; j++)
      if (j != i) {
        const double f = (x - nd[j]) / (nd[i] - nd[j]), df = 1.0 / (nd[i] - nd[j]);
        der = der * f + val * df;
        val *= f;
      }
    dsh[i] = der;
  }
}

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
  } else if (ntd > 3) {
    // cutint/spacetimecutrule.cpp:46-95: 50 subdivisions, exact zeros at the subdivision points are roots, every sign change is
    // refined by at most 70 rounds of (secant point with |value| < 5e-16 ends; else bisection).  A polynomial of degree ntd - 1
    // has at most ntd - 1 zeros, so at most ntd - 1 sign changes; exact zeros at grid points are capped likewise.
    double nd[ST_MAXTD], sh[ST_MAXTD];
    st_nodes(ntd, nd);
    auto eval = [&](double xi) {
      st_lagrange(ntd, nd, xi, sh);
      double v = 0.0;
      for (int k = 0; k < ntd; k++) v += li[k] * sh[k];
      return v;
    };
    constexpr int subdivs = 50;
    const double delta_x = 1. / subdivs;
    double ia[ST_MAXTD], prev = 0.0;
    int ni = 0;
    for (int i = 0; i < subdivs + 1; i++) {
      const double xi = delta_x * i, v = eval(xi);
      if (v == 0 && n < ST_MAXTD - 1) roots[n++] = xi;
      if (i >= 1 && prev * v < 0 && ni < ST_MAXTD - 1) ia[ni++] = xi - delta_x;
      prev = v;
    }
    for (int m = 0; m < ni && n < ST_MAXTD - 1; m++) {
      double a = ia[m], b = ia[m] + delta_x;
      double aval = eval(a), bval = eval(b);
      for (int j = 0; j < 70; j++) {
        const double x_lin = a - aval * (b - a) / (bval - aval);
        double val = eval(x_lin);
        if (2 * fabs(val) < 1e-15) { a = b = x_lin; break; }
        const double x_mid = 0.5 * (a + b);
        val = eval(x_mid);
        if (val == 0) break;
        if (val * aval < 0) { b = x_mid; bval = val; }
        else if (val * bval < 0) { a = x_mid; aval = val; }
        else break;                                    // the reference throws ("Strange sign structure"): cannot happen for a sign change
      }
      roots[n++] = 0.5 * (a + b);
    }
  }
  return n;
}

__device__ __forceinline__ void st_time_shape(int ntd, double t, double *sh) {
  if (ntd == 1) { sh[0] = 1.0; }
  else if (ntd == 2) { sh[0] = 1.0 - t; sh[1] = t; }
  else if (ntd == 3) { sh[0] = (1 - t) * (1 - 2 * t); sh[1] = 4 * t * (1 - t); sh[2] = t * (2 * t - 1); }
  else {
    double nd[ST_MAXTD];
    st_nodes(ntd, nd);
    st_lagrange(ntd, nd, t, sh);
  }
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
  double lst[ST_MAXTD][NV], p[NV][DIM];
#pragma unroll
  for (int k = 0; k < NV; k++) {
    v[k] = __ldg(A.M.el2v + e * NV + k);
    if (POINTS) {
#pragma unroll
      for (int d = 0; d < DIM; d++) p[k][d] = __ldg(A.M.vert + (int64_t)v[k] * DIM + d);
    }
  }
  for (int k = 0; k < A.ntd; k++)
#pragma unroll
    for (int i = 0; i < NV; i++) lst[k][i] = __ldg(A.lset_st + (int64_t)k * A.M.nv + v[i]);
  double cp[ST_MAXCP];
  int ncp = 0;
  cp[ncp++] = 0.0; cp[ncp++] = 1.0;
#pragma unroll
  for (int i = 0; i < NV; i++) {
    double li[ST_MAXTD], r[ST_MAXTD];
    for (int k = 0; k < A.ntd; k++) li[k] = lst[k][i];
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
      double sh[ST_MAXTD], ls[4];
      st_time_shape(A.ntd, t, sh);
#pragma unroll
      for (int k = 0; k < NV; k++) {
        double d = lst[0][k] * sh[0];
        for (int m = 1; m < A.ntd; m++) d += lst[m][k] * sh[m];
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

// ---------------------------------------------------------------------------------------------------------------
// space-time element matrices (SpaceTimeFE, spacetime/SpaceTimeFE.cpp:29-113, inside T_CalcElementMatrixAdd with the space-time
// rule): shape (j, i) = tau_j(tref) phi_i(x), dof j * nd_space + i; block[(j,i)][(j',i')] =
//   sum_q |det J| w_q ( c_m phi_i phi_i' tau_j tau_j' + c_s grad phi_i . grad phi_i' tau_j tau_j' + c_t phi_i tau_j phi_i' tau_j'' )
// (c_t: dt(u) v, the derivative with respect to the REFERENCE time on the trial function, CalcDtShape).  One thread per list
// entry.  All points of one time Gauss point share tau: the thread sums the spatial mass / stiffness matrices (upper triangle, in
// registers / local memory) over them and, when the time changes, adds the time outer products into its own block of the output.
// ---------------------------------------------------------------------------------------------------------------
struct StForm { double cm, cs, ct; int ntd_fe; };

// NTDR > 0: the number of time dofs is the compile-time constant NTDR and the whole (NTDR NS)^2 block lives in registers (P1 in space x
// time order 1: 36 / 64 doubles) -- the time outer products of a flush then cost no memory traffic; NTDR = 0: any time order, the flush
// adds into the thread's block in global memory.
template <int DIM, int ORDER, int NTDR = 0>
struct StMatF {
  static constexpr int NS = H1<DIM, ORDER>::ND, NSYM = NS * (NS + 1) / 2;
  static constexpr int NREG = NTDR > 0 ? NTDR * NS * NTDR * NS : 1;
  double blk[NREG];
  const double (*gl)[DIM];
  double *out;                       // this entry's block, N x N, N = ntd_fe * NS
  StForm f;
  double meas;
  double tnodes[ST_MAXTD];
  double Sm[NSYM], Ss[NSYM];
  double tcur;
  bool have;
  __device__ __forceinline__ void flush() {
    if (!have) return;
    double tau[ST_MAXTD], dtau[ST_MAXTD];
    if (f.ntd_fe == 1) { tau[0] = 1.0; dtau[0] = 0.0; }
    else { st_lagrange(f.ntd_fe, tnodes, tcur, tau); st_dlagrange(f.ntd_fe, tnodes, tcur, dtau); }
    if constexpr (NTDR > 0) {
      constexpr int N = NTDR * NS;
#pragma unroll
      for (int j = 0; j < NTDR; j++)
#pragma unroll
        for (int jp = 0; jp < NTDR; jp++) {
          const double a = tau[j] * tau[jp], b = f.ct * tau[j] * dtau[jp];
          int t = 0;
#pragma unroll
          for (int i = 0; i < NS; i++)
#pragma unroll
            for (int ip = i; ip < NS; ip++, t++) {
              const double v = meas * (a * (f.cm * Sm[t] + f.cs * Ss[t]) + b * Sm[t]);
              blk[(j * NS + i) * N + jp * NS + ip] += v;
              if (ip != i) blk[(j * NS + ip) * N + jp * NS + i] += v;
            }
        }
    } else {
      const int N = f.ntd_fe * NS;
      for (int j = 0; j < f.ntd_fe; j++)
        for (int jp = 0; jp < f.ntd_fe; jp++) {
          const double a = tau[j] * tau[jp], b = f.ct * tau[j] * dtau[jp];
          int t = 0;
          for (int i = 0; i < NS; i++)
            for (int ip = i; ip < NS; ip++, t++) {
              const double v = meas * (a * (f.cm * Sm[t] + f.cs * Ss[t]) + b * Sm[t]);
              out[(j * NS + i) * N + jp * NS + ip] += v;
              if (ip != i) out[(j * NS + ip) * N + jp * NS + i] += v;
            }
        }
    }
#pragma unroll
    for (int t = 0; t < NSYM; t++) Sm[t] = Ss[t] = 0.0;
    have = false;
  }
  __device__ __forceinline__ void operator()(const double *xi, double t, double w) {
    if (have && t != tcur) flush();
    tcur = t; have = true;
    double lam[4] = {xi[0], xi[1], xi[2], 0.0};
    double last = 1.0;
#pragma unroll
    for (int k = 0; k < DIM; k++) last -= lam[k];
    lam[DIM] = last;
    double sv[NS], gv[NS][DIM];
#pragma unroll
    for (int i = 0; i < NS; i++) h1_shape_grad<DIM, ORDER>(i, lam, gl, sv[i], gv[i]);
    int tt = 0;
#pragma unroll
    for (int i = 0; i < NS; i++)
#pragma unroll
      for (int ip = i; ip < NS; ip++, tt++) {
        double gg = 0.0;
#pragma unroll
        for (int d = 0; d < DIM; d++) gg += gv[i][d] * gv[ip][d];
        Sm[tt] += w * sv[i] * sv[ip];
        Ss[tt] += w * gg;
      }
  }
};

template <int DIM, int ORDER, int NTDR>
__global__ void __launch_bounds__(128) k_st_elmat(StArgs A, StForm f, int64_t nlist, const int32_t *__restrict__ els, double *__restrict__ out) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nlist) return;
  const int64_t e = els ? els[k] : k;
  constexpr int NS = H1<DIM, ORDER>::ND;
  const int N = f.ntd_fe * NS;
  double *blk = out + k * (int64_t)N * N;
  if constexpr (NTDR == 0)
    for (int t = 0; t < N * N; t++) blk[t] = 0.0;
  double p[DIM + 1][DIM];
#pragma unroll
  for (int m = 0; m <= DIM; m++) {
    const int v = __ldg(A.M.el2v + e * (DIM + 1) + m);
#pragma unroll
    for (int d = 0; d < DIM; d++) p[m][d] = __ldg(A.M.vert + (int64_t)v * DIM + d);
  }
  ElGeom<DIM> G;
  element_geometry<DIM>(p, G);
  double gl[DIM + 1][DIM];
#pragma unroll
  for (int d = 0; d < DIM; d++) {
    double sgl = 0.0;
#pragma unroll
    for (int m = 0; m < DIM; m++) { gl[m][d] = G.Jinv[m][d]; sgl += G.Jinv[m][d]; }
    gl[DIM][d] = -sgl;
  }
  StMatF<DIM, ORDER, NTDR> F;
  F.gl = gl; F.out = blk; F.f = f; F.meas = fabs(G.det); F.have = false; F.tcur = 0.0;
  st_nodes(f.ntd_fe, F.tnodes);
#pragma unroll
  for (int t = 0; t < StMatF<DIM, ORDER, NTDR>::NSYM; t++) F.Sm[t] = F.Ss[t] = 0.0;
#pragma unroll
  for (int t = 0; t < StMatF<DIM, ORDER, NTDR>::NREG; t++) F.blk[t] = 0.0;
  st_walk<DIM, true>(A, e, F);
  F.flush();
  if constexpr (NTDR > 0) {
#pragma unroll
    for (int t = 0; t < StMatF<DIM, ORDER, NTDR>::NREG; t++) blk[t] = F.blk[t];
  }
}

// space-time CutInformation::Update (xfem/cutinfo.cpp:79-111 with time_order >= 0): part "volumes" of the space-time prism from the
// NEG and POS rules of order 0 in space -> domain type and cut ratio per element
struct SumF {
  double s = 0.0;
  __device__ __forceinline__ void operator()(const double *, double, double w) { s += w; }
};

template <int DIM>
__global__ void __launch_bounds__(128) k_st_classify(StArgs A, int8_t *__restrict__ dom, double *__restrict__ ratio) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= A.M.ne) return;
  double pv[2];
  for (int np = 1; np >= 0; np--) {                  // POS first, as the reference
    StArgs B = A;
    B.dt = np;
    SumF F;
    st_walk<DIM, true>(B, e, F);
    pv[np] = F.s;
  }
  ratio[e] = pv[NEG] / (pv[NEG] + pv[POS]);
  dom[e] = (int8_t)(pv[NEG] > 0.0 ? (pv[POS] > 0.0 ? IF : NEG) : POS);
}

static int st_args(const ngsx_mesh *mesh, const double *lset_st, int ntd, int dt, int order_space, int order_time, StArgs &A) {
  NGSX_REQUIRE(mesh && (mesh->dim == 2 || mesh->dim == 3) && mesh->vert && mesh->el2v && lset_st, "space-time rule: bad mesh / level set");
  NGSX_REQUIRE(dt >= 0 && dt <= 2, "space-time rule: domain_type must be NEG, POS or IF");
  if (ntd < 1 || ntd > ST_MAXTD) { set_error("space-time rule: time element of order > 5 (NodalTimeFE has no nodes for it either)"); return NGSX_ERR_UNSUPPORTED; }
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

int ngsx_st_elmat_h1(const ngsx_mesh *mesh, const double *lset_st, int ntime_dofs, int fe_order, int fe_time_dofs, int dt, int order_space,
                     int order_time, double c_mass, double c_stiff, double c_dt, int64_t nlist, const int32_t *els, double *elmats, void *stream) {
  StArgs A{};
  if (int rc = st_args(mesh, lset_st, ntime_dofs, dt, order_space, order_time, A)) return rc;
  NGSX_REQUIRE(nlist >= 0 && (nlist == 0 || elmats), "st_elmat_h1: bad arguments");
  if (fe_order < 1 || fe_order > 2 || fe_time_dofs < 1 || fe_time_dofs > ST_MAXTD) {
    set_error("st_elmat_h1: H1 order 1..2 in space and a nodal time element of order 0..5 are on the CUDA path");
    return NGSX_ERR_UNSUPPORTED;
  }
  if (nlist == 0) return NGSX_OK;
  StForm f{c_mass, c_stiff, c_dt, fe_time_dofs};
  const unsigned nb = (unsigned)((nlist + 127) / 128);
  cudaStream_t st = as_stream(stream);
  switch (mesh->dim * 10 + fe_order) {
    case 21:
      if (fe_time_dofs == 2) k_st_elmat<2, 1, 2><<<nb, 128, 0, st>>>(A, f, nlist, els, elmats);      // register-resident 6 x 6 block
      else k_st_elmat<2, 1, 0><<<nb, 128, 0, st>>>(A, f, nlist, els, elmats);
      break;
    case 22: k_st_elmat<2, 2, 0><<<nb, 128, 0, st>>>(A, f, nlist, els, elmats); break;
    case 31:
      if (fe_time_dofs == 2) k_st_elmat<3, 1, 2><<<nb, 128, 0, st>>>(A, f, nlist, els, elmats);      // 8 x 8
      else k_st_elmat<3, 1, 0><<<nb, 128, 0, st>>>(A, f, nlist, els, elmats);
      break;
    case 32: k_st_elmat<3, 2, 0><<<nb, 128, 0, st>>>(A, f, nlist, els, elmats); break;
  }
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_st_classify(const ngsx_mesh *mesh, const double *lset_st, int ntime_dofs, int time_order, int8_t *dom, double *ratio, void *stream) {
  StArgs A{};
  if (int rc = st_args(mesh, lset_st, ntime_dofs, NEG, 0, time_order, A)) return rc;
  if (mesh->ne == 0) return NGSX_OK;
  NGSX_REQUIRE(dom && ratio, "st_classify: NULL array");
  const unsigned nb = (unsigned)((mesh->ne + 127) / 128);
  if (mesh->dim == 2) k_st_classify<2><<<nb, 128, 0, as_stream(stream)>>>(A, dom, ratio);
  else k_st_classify<3><<<nb, 128, 0, as_stream(stream)>>>(A, dom, ratio);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

}  // extern "C"
