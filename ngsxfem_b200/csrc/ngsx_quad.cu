// ngsx_quad.cu -- SURVEY section 8 row f-5: straight-cut quadrature on QUADRILATERALS and HEXAHEDRA.
//
// Reference behaviour (what, not how): LevelsetCutQuadrilateral (cutint/straightcutrule.cpp:206-501) under the default policy
// FIND_OPTIMAL, and the ET_QUAD / ET_HEX branch of StraightCutIntegrationRule[Untransformed] (:580-726).  The level set is the
// multilinear interpolant of the element's vertex values (LevelsetWrapper, :503-556).  A cut cell is integrated as a tensor rule:
// Gauss points along a direction xi, and at every height the rule of the (D-1)-dimensional cut of the slice -- a cut segment for
// quadrilaterals, a cut quadrilateral for hexahedra.  The direction is the one along which the zero level is a graph with the
// flattest slope (exact criterion in 2D :392-434, sufficient criterion in 3D :353-390, 436-451; the cell is mirrored into the
// chosen direction, :305-351); where the sign of the level set changes along a xi-edge the cell is first split at those heights
// (:206-251); if no direction qualifies the cell falls back to simplices (:470-489).
//
// Execution model: ONE THREAD PER ELEMENT walks its rule and hands every point to a functor -- count, emit at scan offsets in
// the reference's order, or integrate (as the space-time walker, ngsx_spacetime.cu).  The branchy case analysis stays inside
// the thread; uncut cells (the vast majority of a mesh) take the short tensor-rule path.  Pinned against the rules of the
// compiled reference (tests/golden/quad_hex_rules.npz) and oracle/quad_ref.py.
#include <algorithm>
#include "ngsx_host.cuh"

namespace ngsx {

struct QRules {
  const double4 *r1; int n1;      // Gauss rule on the segment (SelectIntegrationRule(ET_SEGM, order))
  const double4 *r2; int n2;      // triangle rule (2D fallback, interface pieces of the 3D fallback)
  const CutTable *tab;
};

__device__ __forceinline__ int q_check(const double *v, int n, double eps) {      // CheckIfStraightCut (:11-39)
  bool haspos = false, hasneg = false;
  for (int i = 0; i < n; i++) { haspos |= v[i] > eps; hasneg |= v[i] < -eps; }
  if (haspos == hasneg) return IF;
  return hasneg ? NEG : POS;
}

// bilinear level set on a quadrilateral: c00 + c10 x + c01 y + c11 x y (GetCoeffsFromVals, ET_QUAD)
struct Bil {
  double c00, c10, c01, c11;
  __device__ __forceinline__ double operator()(double x, double y) const { return c00 + c10 * x + c01 * y + c11 * x * y; }
  __device__ __forceinline__ double gx(double y) const { return c10 + c11 * y; }
  __device__ __forceinline__ double gy(double x) const { return c01 + c11 * x; }
};
__device__ __forceinline__ Bil bil_from_vals(const double *v) {
  Bil c;
  c.c00 = v[0]; c.c10 = v[1] - c.c00; c.c01 = v[3] - c.c00; c.c11 = v[2] - c.c10 - c.c01 - c.c00;
  return c;
}

// tensor rule of a sub-rectangle (Quadrilateral::GetPlainIntegrationRule, :102-128): x outer, y inner
template <class F>
__device__ __forceinline__ void q_plain(double x0, double x1, double y0, double y1, const QRules &R, F &f) {
  const double vol = fabs((x1 - x0) * (y1 - y0));
  for (int i = 0; i < R.n1; i++)
    for (int j = 0; j < R.n1; j++)
      f(x0 + (x1 - x0) * R.r1[i].x, y0 + (y1 - y0) * R.r1[j].x, (R.r1[i].w * R.r1[j].w) * vol);
}

// GetTensorProductAlongXiIntegrationRule (:255-303), D = 2, xi = y: per Gauss height the rule of the cut segment (vertex 0 of
// ET_SEGM sits at xhat = 1).  false: the reference throws (a slice that is not cut, or a vertex value exactly zero)
template <class F>
__device__ bool q_tensor(const Bil &c, int dt, double x0, double x1, double y0, double y1, const QRules &R, F &f) {
  for (int i = 0; i < R.n1; i++) {
    const double ya = y0 + R.r1[i].x * (y1 - y0), w1 = R.r1[i].w;
    const double pr[2] = {c(x1, ya), c(x0, ya)};
    if (q_check(pr, 2, 0.0) != IF || pr[0] == 0.0 || pr[1] == 0.0) return false;
    const double cp = 1.0 + pr[0] / (pr[0] - pr[1]) * (0.0 - 1.0);               // CalcIFPolytopEUsingLset on the line (points 1, 0)
    if (dt == IF) {
      const double x = x0 + cp * (x1 - x0);
      const double g0 = c.gx(ya), g1 = c.gy(x);
      f(x, ya, 1.0 * w1 * (y1 - y0) * (sqrt(g0 * g0 + g1 * g1) / fabs(g0)));
    } else {
      const bool rel0 = dt == POS ? pr[0] >= 0.0 : pr[0] < 0.0;                   // the one vertex on the dt side
      const double v = rel0 ? 1.0 : 0.0, len = fabs(v - cp);
      for (int q = 0; q < R.n1; q++) {
        const double xi = R.r1[q].x;
        const double xh = (1.0 - xi) * cp + xi * v;
        f(x0 + xh * (x1 - x0), ya, (R.r1[q].w * len) * w1 * (y1 - y0));
      }
    }
  }
  return true;
}

// GetIntegrationRuleAlongXi (:453-468) with HasTopologyChangeAlongXi / Decompose (:206-251)
template <class F>
__device__ bool q_along(const Bil &c, int dt, const QRules &R, F &f) {
  const double cv[4] = {c(0.0, 0.0), c(1.0, 0.0), c(1.0, 1.0), c(0.0, 1.0)};
  const int d = q_check(cv, 4, 0.0);
  if (d != IF) {
    if (d == dt) q_plain(0.0, 1.0, 0.0, 1.0, R, f);
    return true;
  }
  double xs[4] = {0.0, 1.0, 0.0, 0.0};
  int nx = 2;
  const int ea[2] = {1, 0}, eb[2] = {2, 3};
  for (int k = 0; k < 2; k++) {
    const double v[2] = {cv[eb[k]], cv[ea[k]]};
    if (q_check(v, 2, 0.0) == IF) xs[nx++] = 1.0 + v[0] / (v[0] - v[1]) * (0.0 - 1.0);
  }
  if (nx == 2) return q_tensor(c, dt, 0.0, 1.0, 0.0, 1.0, R, f);
  for (int a = 1; a < nx; a++) {                                   // sorted set of heights
    const double v = xs[a];
    int b = a;
    while (b > 0 && xs[b - 1] > v) { xs[b] = xs[b - 1]; b--; }
    xs[b] = v;
  }
  int nu = 1;
  for (int a = 1; a < nx; a++) if (xs[a] != xs[nu - 1]) xs[nu++] = xs[a];
  for (int k = 0; k + 1 < nu; k++) {
    const double y0 = xs[k], y1 = xs[k + 1];
    const double sv[4] = {c(0.0, y0), c(1.0, y0), c(1.0, y1), c(0.0, y1)};
    const int d2 = q_check(sv, 4, 1e-15);
    if (d2 == IF) { if (!q_tensor(c, dt, 0.0, 1.0, y0, y1, R, f)) return false; }
    else if (d2 == dt) q_plain(0.0, 1.0, y0, y1, R, f);
  }
  return true;
}

// cut rule of a general triangle whose level set is the P1 interpolant of its (clamped) vertex values: the pieces of the 2D
// fallback (GetFallbackIntegrationRule, :470-489).  P: vertices in the coordinates of the quadrilateral.
template <class F>
__device__ bool q_triangle(const double (*P)[2], const double *vals, int dt, const QRules &R, F &f) {
  const int d = q_check(vals, 3, 0.0);
  if (d != IF) {
    if (d != dt) return true;
    const double det = fabs((P[1][0] - P[0][0]) * (P[2][1] - P[0][1]) - (P[2][0] - P[0][0]) * (P[1][1] - P[0][1]));
    for (int q = 0; q < R.n2; q++) {                                // SimpleX::GetPlainIntegrationRule: base point P0
      const double a = R.r2[q].x, b = R.r2[q].y, o = 1.0 - a - b;
      f(o * P[0][0] + a * P[1][0] + b * P[2][0], o * P[0][1] + a * P[1][1] + b * P[2][1], R.r2[q].w * det);
    }
    return true;
  }
  int mask;
  if (sign_test<2>(vals, mask) != IF) return false;                 // a vertex value exactly zero on a cut triangle: the reference throws
  if (dt != IF) {
    const double det = fabs((P[0][0] - P[2][0]) * (P[1][1] - P[2][1]) - (P[1][0] - P[2][0]) * (P[0][1] - P[2][1]));
    const CutEntry E = R.tab->e[0][dt][mask];
    for (int s = 0; s < E.ns; s++) {
      double S[4][3];
      for (int a = 0; a < E.npt; a++) coded_point<2>(E.v[s][a], vals, S[a]);
      const double fac = simplex_factor(E.npt, S);
      for (int q = 0; q < R.n2; q++) {
        double xi[3];
        simplex_point(E.npt, S, R.r2[q], xi);
        const double o = 1.0 - xi[0] - xi[1];
        f(xi[0] * P[0][0] + xi[1] * P[1][0] + o * P[2][0], xi[0] * P[0][1] + xi[1] * P[1][1] + o * P[2][1], (R.r2[q].w * fac) * det);
      }
    }
    return true;
  }
  double cut[2][2];
  int nc = 0;
  for (int i = 0; i < 3; i++)
    for (int j = i + 1; j < 3; j++)
      if ((vals[i] >= 0.0) != (vals[j] >= 0.0)) {
        if (nc == 2) return false;
        const double t = vals[i] / (vals[i] - vals[j]);
        cut[nc][0] = P[i][0] + t * (P[j][0] - P[i][0]);
        cut[nc][1] = P[i][1] + t * (P[j][1] - P[i][1]);
        nc++;
      }
  if (nc != 2) return false;
  const double dx = cut[1][0] - cut[0][0], dy = cut[1][1] - cut[0][1], len = sqrt(dx * dx + dy * dy);
  for (int q = 0; q < R.n1; q++) {
    const double s = R.r1[q].x;
    f((1.0 - s) * cut[0][0] + s * cut[1][0], (1.0 - s) * cut[0][1] + s * cut[1][1], R.r1[q].w * len);
  }
  return true;
}

// LevelsetCutQuadrilateral::GetIntegrationRule on the unit square, policy FIND_OPTIMAL (:392-434, 491-500).
// f(x, y, w) in the coordinates of the square.  false: the reference throws.
template <class F>
__device__ bool q_cut_quad(const double *vals, int dt, const QRules &R, F &f) {
  const Bil c = bil_from_vals(vals);
  // GetExactCritsQBound2D (:392-413): IEEE semantics of the reference (division by c11 = 0 gives inf / nan, comparisons false)
  bool allow[2] = {true, true};
  double h = -c.c10 / c.c11;
  if (0.0 < h && h < 1.0) allow[1] = false;
  h = -c.c01 / c.c11;
  if (0.0 < h && h < 1.0) allow[0] = false;
  double qmax[2] = {0.0, 0.0};
  const double px[4] = {0.0, 1.0, 1.0, 0.0}, py[4] = {0.0, 0.0, 1.0, 1.0};
  for (int k = 0; k < 4; k++) {
    const double g0 = c.gx(py[k]), g1 = c.gy(px[k]);
    const double nrm = sqrt(g0 * g0 + g1 * g1);
    const double qy = fabs(g1) / nrm, qx = fabs(g0) / nrm;
    if (qy > qmax[1]) qmax[1] = qy;
    if (qx > qmax[0]) qmax[0] = qx;
  }
  if (!allow[0]) qmax[0] = 2.0;
  if (!allow[1]) qmax[1] = 2.0;
  const double CB = 0.999;
  int sw;                                                           // 0: along y as given, 1: x and y swapped, 2: fallback
  if (qmax[0] < CB && qmax[1] < CB) sw = qmax[1] <= qmax[0] ? 0 : 1;
  else if (qmax[1] < CB) sw = 0;
  else if (qmax[0] < CB) sw = 1;
  else sw = 2;
  if (sw == 0) return q_along(c, dt, R, f);
  if (sw == 1) {                                                    // GetIntegrationRuleOnXYPermutatedQuad (:305-320)
    const Bil cr{c.c00, c.c01, c.c10, c.c11};
    auto g = [&](double x, double y, double w) { f(y, x, w); };
    return q_along(cr, dt, R, g);
  }
  const int tri[2][3] = {{0, 1, 3}, {2, 1, 3}};
  for (int t = 0; t < 2; t++) {
    double P[3][2], tv[3];
    for (int a = 0; a < 3; a++) {
      P[a][0] = px[tri[t][a]]; P[a][1] = py[tri[t][a]];
      const double dval = c(P[a][0], P[a][1]);
      tv[a] = fabs(dval) > 1e-14 ? dval : 1e-14;                    // LevelsetWrapper::update_initial_coefs (:547-556)
    }
    if (!q_triangle(P, tv, dt, R, f)) return false;
  }
  return true;
}

// ---- hexahedra ------------------------------------------------------------------------------------------------------
// trilinear level set  sum c[i][j][k] x^i y^j z^k  (GetCoeffsFromVals, ET_HEX :513-525)
struct Tril {
  double c[2][2][2];
  __device__ __forceinline__ double operator()(double x, double y, double z) const {
    return c[0][0][0] + c[1][0][0] * x + c[0][1][0] * y + c[1][1][0] * x * y +
           (c[0][0][1] + c[1][0][1] * x + c[0][1][1] * y + c[1][1][1] * x * y) * z;
  }
  __device__ __forceinline__ void grad(double x, double y, double z, double *g) const {      // GetGrad (:527-533)
    g[0] = c[1][0][0] + c[1][1][0] * y + c[1][0][1] * z + c[1][1][1] * y * z;
    g[1] = c[0][1][0] + c[1][1][0] * x + c[0][1][1] * z + c[1][1][1] * x * z;
    g[2] = c[0][0][1] + c[0][1][1] * y + c[1][0][1] * x + c[1][1][1] * x * y;
  }
};
__device__ __forceinline__ Tril tril_from_vals(const double *v) {
  Tril t;
  t.c[0][0][0] = v[0]; t.c[1][0][0] = v[1] - v[0]; t.c[0][1][0] = v[3] - v[0]; t.c[0][0][1] = v[4] - v[0];
  t.c[1][1][0] = v[2] - t.c[1][0][0] - t.c[0][1][0] - t.c[0][0][0];
  t.c[1][0][1] = v[5] - t.c[1][0][0] - t.c[0][0][1] - t.c[0][0][0];
  t.c[0][1][1] = v[7] - t.c[0][1][0] - t.c[0][0][1] - t.c[0][0][0];
  t.c[1][1][1] = v[6] - t.c[1][1][0] - t.c[1][0][1] - t.c[0][1][1] - t.c[1][0][0] - t.c[0][1][0] - t.c[0][0][1] - t.c[0][0][0];
  return t;
}

template <class F>
__device__ __forceinline__ void h_plain(double z0, double z1, const QRules &R, F &f) {
  const double vol = fabs(z1 - z0);
  for (int i = 0; i < R.n1; i++)
    for (int j = 0; j < R.n1; j++)
      for (int k = 0; k < R.n1; k++)
        f(R.r1[i].x, R.r1[j].x, z0 + (z1 - z0) * R.r1[k].x, (R.r1[i].w * R.r1[j].w * R.r1[k].w) * vol);
}

// tensor rule along z on the slab [z0, z1] of the unit cube: the slices are cut quadrilaterals (child LevelsetCutQuadrilateral)
template <class F>
__device__ bool h_tensor(const Tril &c, int dt, double z0, double z1, const QRules &R, F &f) {
  for (int i = 0; i < R.n1; i++) {
    const double za = z0 + R.r1[i].x * (z1 - z0), w1 = R.r1[i].w;
    const double pr[4] = {c(0.0, 0.0, za), c(1.0, 0.0, za), c(1.0, 1.0, za), c(0.0, 1.0, za)};
    auto g = [&](double x, double y, double w) {
      double scale = 1.0;
      if (dt == IF) {
        double gr[3];
        c.grad(x, y, za, gr);
        scale = sqrt(gr[0] * gr[0] + gr[1] * gr[1] + gr[2] * gr[2]) / sqrt(gr[0] * gr[0] + gr[1] * gr[1]);
      }
      f(x, y, za, w * w1 * (z1 - z0) * scale);
    };
    if (!q_cut_quad(pr, dt, R, g)) return false;
  }
  return true;
}

template <class F>
__device__ bool h_along(const Tril &c, int dt, const QRules &R, F &f) {
  const double px[8] = {0, 1, 1, 0, 0, 1, 1, 0}, py[8] = {0, 0, 1, 1, 0, 0, 1, 1}, pz[8] = {0, 0, 0, 0, 1, 1, 1, 1};
  double cv[8];
  for (int k = 0; k < 8; k++) cv[k] = c(px[k], py[k], pz[k]);
  const int d = q_check(cv, 8, 0.0);
  if (d != IF) {
    if (d == dt) h_plain(0.0, 1.0, R, f);
    return true;
  }
  double xs[6] = {0.0, 1.0, 0.0, 0.0, 0.0, 0.0};
  int nx = 2;
  for (int a = 0; a < 4; a++) {                                     // the four z-edges (a, a + 4)
    const double v[2] = {cv[a + 4], cv[a]};
    if (q_check(v, 2, 0.0) == IF) xs[nx++] = 1.0 + v[0] / (v[0] - v[1]) * (0.0 - 1.0);
  }
  if (nx == 2) return h_tensor(c, dt, 0.0, 1.0, R, f);
  for (int a = 1; a < nx; a++) {
    const double v = xs[a];
    int b = a;
    while (b > 0 && xs[b - 1] > v) { xs[b] = xs[b - 1]; b--; }
    xs[b] = v;
  }
  int nu = 1;
  for (int a = 1; a < nx; a++) if (xs[a] != xs[nu - 1]) xs[nu++] = xs[a];
  for (int k = 0; k + 1 < nu; k++) {
    const double z0 = xs[k], z1 = xs[k + 1];
    double sv[8];
    for (int m = 0; m < 8; m++) sv[m] = c(px[m], py[m], m < 4 ? z0 : z1);
    const int d2 = q_check(sv, 8, 1e-15);
    if (d2 == IF) { if (!h_tensor(c, dt, z0, z1, R, f)) return false; }
    else if (d2 == dt) h_plain(z0, z1, R, f);
  }
  return true;
}

// LevelsetCutQuadrilateral::GetIntegrationRule, D = 3, FIND_OPTIMAL (:353-390, 436-451).  status: 0 ok, 1 the reference throws,
// 2 the fallback to six tetrahedra would be taken (:470-489; needs, for every direction d, a corner with g_d^2 >= 0.998 of the
// sum of the per-direction maxima: not reachable with a finite gradient, not on the CUDA path)
template <class F>
__device__ int h_cut_hex(const double *vals, int dt, const QRules &R, F &f) {
  const Tril c = tril_from_vals(vals);
  double g[8][3], mx[3] = {0.0, 0.0, 0.0};
  for (int k = 0; k < 8; k++) {
    c.grad((double)(k & 1), (double)((k >> 1) & 1), (double)(k >> 2), g[k]);
    for (int d = 0; d < 3; d++) mx[d] = fmax(mx[d], g[k][d] * g[k][d]);
  }
  const double V = sqrt(mx[0] + mx[1] + mx[2]);                      // GetSufficientCritsQBound
  double sb[3];
  bool isnan_any = false;
  for (int d = 0; d < 3; d++) {
    double qmax = 0.0;
    for (int k = 0; k < 8; k++) {
      const double est = (V * V) / (V * V - g[k][d] * g[k][d]);
      if (est > qmax) qmax = est;
    }
    sb[d] = sqrt(1.0 - 1.0 / qmax);
    isnan_any |= sb[d] != sb[d];
  }
  if (isnan_any) return 1;
  int md = 0;
  for (int d = 1; d < 3; d++) if (sb[d] <= sb[md]) md = d;           // min_element with a `<=` comparator: the last of equal minima
  if (!(sb[md] < 0.999)) return 2;
  if (md == 2) return h_along(c, dt, R, f) ? 0 : 1;
  Tril cr;
  for (int i = 0; i < 2; i++)
    for (int j = 0; j < 2; j++)
      for (int k = 0; k < 2; k++) cr.c[i][j][k] = md == 0 ? c.c[k][j][i] : c.c[i][k][j];      // X_Z (:322-336) / Y_Z (:338-351)
  if (md == 0) {
    auto h = [&](double x, double y, double z, double w) { f(z, y, x, w); };
    return h_along(cr, dt, R, h) ? 0 : 1;
  }
  auto h = [&](double x, double y, double z, double w) { f(x, z, y, w); };
  return h_along(cr, dt, R, h) ? 0 : 1;
}

// StraightCutIntegrationRuleUntransformed for ET_QUAD / ET_HEX (:663-726): f(x, y, z, w) over the rule of element values `vals`.
// returns 0 ok (also: no rule), 1 throws, 2 unsupported fallback
template <int DIM, class F>
__device__ int cell_walk(const double *vals, int dt, const QRules &R, F &f) {
  constexpr int NV = 1 << DIM;
  const int d = q_check(vals, NV, 0.0);
  if (d != IF) {
    if (d != dt) return 0;
    if constexpr (DIM == 2) { auto g = [&](double x, double y, double w) { f(x, y, 0.0, w); }; q_plain(0.0, 1.0, 0.0, 1.0, R, g); }
    else h_plain(0.0, 1.0, R, f);
    return 0;
  }
  if constexpr (DIM == 2) {
    auto g = [&](double x, double y, double w) { f(x, y, 0.0, w); };
    return q_cut_quad(vals, dt, R, g) ? 0 : 1;
  } else {
    return h_cut_hex(vals, dt, R, f);
  }
}

// ---- element geometry: multilinear map of the 2^DIM vertices (ScalarFE<ET_QUAD / ET_HEX, 1>, NGS/fem/elementtransformation.cpp:68-87)
template <int DIM> struct CellGeom {
  static constexpr int NV = 1 << DIM;
  double p[NV][DIM];
  __device__ __forceinline__ void load(const ngsx_mesh &M, int64_t e, double *vals, const double *__restrict__ lset) {
#pragma unroll
    for (int k = 0; k < NV; k++) {
      const int v = __ldg(M.el2v + e * NV + k);
      vals[k] = __ldg(lset + v);
#pragma unroll
      for (int d = 0; d < DIM; d++) p[k][d] = __ldg(M.vert + (int64_t)v * DIM + d);
    }
  }
  // x(xi), Jacobian J[d][k] = dx_d / dxi_k
  __device__ __forceinline__ void map(const double *xi, double *x, double (*J)[DIM]) const {
    double N[NV], dN[NV][DIM];
    const double a = xi[0], b = xi[1];
    if constexpr (DIM == 2) {
      const double n[4] = {(1 - a) * (1 - b), a * (1 - b), a * b, (1 - a) * b};
      const double da[4] = {-(1 - b), (1 - b), b, -b}, db[4] = {-(1 - a), -a, a, (1 - a)};
      for (int k = 0; k < 4; k++) { N[k] = n[k]; dN[k][0] = da[k]; dN[k][1] = db[k]; }
    } else {
      const double c = xi[2];
      const double n2[4] = {(1 - a) * (1 - b), a * (1 - b), a * b, (1 - a) * b};
      const double da[4] = {-(1 - b), (1 - b), b, -b}, db[4] = {-(1 - a), -a, a, (1 - a)};
      for (int k = 0; k < 4; k++) {
        N[k] = n2[k] * (1 - c); N[k + 4] = n2[k] * c;
        dN[k][0] = da[k] * (1 - c); dN[k + 4][0] = da[k] * c;
        dN[k][1] = db[k] * (1 - c); dN[k + 4][1] = db[k] * c;
        dN[k][2] = -n2[k]; dN[k + 4][2] = n2[k];
      }
    }
#pragma unroll
    for (int d = 0; d < DIM; d++) {
      double s = 0.0;
      for (int k = 0; k < NV; k++) s += N[k] * p[k][d];
      x[d] = s;
#pragma unroll
      for (int m = 0; m < DIM; m++) {
        double t = 0.0;
        for (int k = 0; k < NV; k++) t += dN[k][m] * p[k][d];
        J[d][m] = t;
      }
    }
  }
};

template <int DIM>
__device__ __forceinline__ double q_det_inv(const double (*J)[DIM], double (*I)[DIM]) {
  if constexpr (DIM == 2) {
    const double det = J[0][0] * J[1][1] - J[0][1] * J[1][0], id = 1.0 / det;
    I[0][0] = id * J[1][1]; I[0][1] = -id * J[0][1]; I[1][0] = -id * J[1][0]; I[1][1] = id * J[0][0];
    return det;
  } else {
    const double h0 = J[1][1] * J[2][2] - J[1][2] * J[2][1], h1 = J[1][2] * J[2][0] - J[1][0] * J[2][2], h2 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
    const double det = J[0][0] * h0 + J[0][1] * h1 + J[0][2] * h2, id = 1.0 / det;
    I[0][0] = id * h0; I[1][0] = id * h1; I[2][0] = id * h2;
    I[0][1] = id * (J[0][2] * J[2][1] - J[0][1] * J[2][2]); I[1][1] = id * (J[0][0] * J[2][2] - J[0][2] * J[2][0]); I[2][1] = id * (J[0][1] * J[2][0] - J[0][0] * J[2][1]);
    I[0][2] = id * (J[0][1] * J[1][2] - J[0][2] * J[1][1]); I[1][2] = id * (J[0][2] * J[1][0] - J[0][0] * J[1][2]); I[2][2] = id * (J[0][0] * J[1][1] - J[0][1] * J[1][0]);
    return det;
  }
}

// |F^-T n_ref| at a point of the interface rule, n_ref = grad(lset)/|grad(lset)| in reference coordinates (:558-576), and |det F|
template <int DIM>
__device__ __forceinline__ void q_point_factors(const CellGeom<DIM> &G, const double *vals, const double *xi, bool iface, double *x,
                                                double &meas, double &ifscale) {
  double J[DIM][DIM], I[DIM][DIM];
  G.map(xi, x, J);
  meas = fabs(q_det_inv<DIM>(J, I));
  ifscale = 1.0;
  if (iface) {
    double g[3] = {0.0, 0.0, 0.0};
    if constexpr (DIM == 2) { const Bil c = bil_from_vals(vals); g[0] = c.gx(xi[1]); g[1] = c.gy(xi[0]); }
    else { const Tril c = tril_from_vals(vals); c.grad(xi[0], xi[1], xi[2], g); }
    double n2 = 0.0;
#pragma unroll
    for (int d = 0; d < DIM; d++) n2 += g[d] * g[d];
    const double inv = 1.0 / sqrt(n2);
    double s2 = 0.0;
#pragma unroll
    for (int d = 0; d < DIM; d++) {                                 // (F^-T n)_d = sum_k Finv[k][d] n_k
      double t = 0.0;
#pragma unroll
      for (int k = 0; k < DIM; k++) t += I[k][d] * (g[k] * inv);
      s2 += t * t;
    }
    ifscale = sqrt(s2);
  }
}

struct QArgs {
  ngsx_mesh M;
  const double *lset;
  int dt;
  QRules R;
  unsigned long long *err;
};

__device__ __forceinline__ void q_flag(const QArgs &A, int64_t e) {
  if (A.err) {
    atomicAdd(&A.err[0], 1ULL);
    atomicMin(reinterpret_cast<long long *>(&A.err[1]), (long long)e);
  }
}

template <int DIM>
__global__ void __launch_bounds__(128) k_q_count(QArgs A, int32_t *__restrict__ npts) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= A.M.ne) return;
  double vals[1 << DIM];
#pragma unroll
  for (int k = 0; k < (1 << DIM); k++) vals[k] = __ldg(A.lset + __ldg(A.M.el2v + e * (1 << DIM) + k));
  int n = 0;
  auto f = [&](double, double, double, double) { n++; };
  const int st = cell_walk<DIM>(vals, A.dt, A.R, f);
  if (st) { q_flag(A, e); n = 0; }
  npts[e] = n;
}

template <int DIM>
__global__ void __launch_bounds__(128) k_q_fill(QArgs A, int transform_if, const int64_t *__restrict__ offs, double *__restrict__ pts3,
                                               double *__restrict__ wts) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= A.M.ne) return;
  const int64_t o0 = offs[e], cap = offs[e + 1] - o0;
  if (cap == 0) return;
  CellGeom<DIM> G;
  double vals[1 << DIM];
  G.load(A.M, e, vals, A.lset);
  const bool tr = transform_if && A.dt == IF;
  int64_t n = 0;
  auto f = [&](double x, double y, double z, double w) {
    if (n < cap) {
      const double xi[3] = {x, y, z};
      if (tr) {
        double xp[3], meas, sc;
        q_point_factors<DIM>(G, vals, xi, true, xp, meas, sc);
        w *= sc;
      }
      pts3[3 * (o0 + n)] = x; pts3[3 * (o0 + n) + 1] = y; pts3[3 * (o0 + n) + 2] = z;
      wts[o0 + n] = w;
    }
    n++;
  };
  cell_walk<DIM>(vals, A.dt, A.R, f);
}

struct QPoly { int n; double c[NGSX_POLY_MAXTERMS]; int ex[NGSX_POLY_MAXTERMS][3]; };

template <int DIM>
__global__ void __launch_bounds__(128) k_q_integrate(QArgs A, QPoly P, double *__restrict__ partial) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double val = 0.0;
  if (e < A.M.ne) {
    CellGeom<DIM> G;
    double vals[1 << DIM];
    G.load(A.M, e, vals, A.lset);
    const bool iface = A.dt == IF;
    auto f = [&](double x, double y, double z, double w) {
      const double xi[3] = {x, y, z};
      double xp[3] = {0.0, 0.0, 0.0}, meas, sc;
      q_point_factors<DIM>(G, vals, xi, iface, xp, meas, sc);
      double fx = 0.0;
      for (int m = 0; m < P.n; m++) fx += P.c[m] * ipow(xp[0], P.ex[m][0]) * ipow(xp[1], P.ex[m][1]) * ipow(xp[2], P.ex[m][2]);
      val += w * sc * meas * fx;
    };
    if (cell_walk<DIM>(vals, A.dt, A.R, f)) { q_flag(A, e); val = 0.0; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
  if ((threadIdx.x & 31) == 0) partial[e >> 5] = val;
}

// CutInformation::Update on tensor cells (xfem/cutinfo.cpp:54-120): domain from the vertex signs, cut ratio from the order-0 rules
template <int DIM>
__global__ void __launch_bounds__(128) k_q_classify(QArgs A, int8_t *__restrict__ dom, double *__restrict__ ratio) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= A.M.ne) return;
  double vals[1 << DIM];
#pragma unroll
  for (int k = 0; k < (1 << DIM); k++) vals[k] = __ldg(A.lset + __ldg(A.M.el2v + e * (1 << DIM) + k));
  int d = q_check(vals, 1 << DIM, 0.0);
  double r = d == NEG ? 1.0 : 0.0;
  if (d == IF) {
    double vol[2] = {0.0, 0.0};
    int bad = 0;
    for (int side = 0; side < 2; side++) {
      auto f = [&](double, double, double, double w) { vol[side] += w; };
      bad |= cell_walk<DIM>(vals, side, A.R, f);
    }
    if (bad) { q_flag(A, e); d = -1; r = __longlong_as_double(0x7ff8000000000000LL); }
    else r = vol[0] / (vol[0] + vol[1]);
  }
  if (dom) dom[e] = (int8_t)d;
  if (ratio) ratio[e] = r;
}

__global__ void __launch_bounds__(256) k_q_reduce(int64_t n, const double *__restrict__ in, double *__restrict__ partial) {
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
__global__ void __launch_bounds__(1024) k_q_final(int64_t n, const double *__restrict__ partial, double *__restrict__ out) {
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

static int q_args(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, int64_t *err, cudaStream_t st, QArgs &A) {
  NGSX_REQUIRE(mesh && (mesh->dim == 2 || mesh->dim == 3) && lset_v, "quad / hex rule: bad mesh / level set");
  NGSX_REQUIRE(mesh->ne == 0 || (mesh->vert && mesh->el2v), "quad / hex rule: mesh arrays are NULL");
  NGSX_REQUIRE(dt >= 0 && dt <= 2, "quad / hex rule: domain_type must be NEG, POS or IF");
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  A.M = *mesh; A.lset = lset_v; A.dt = dt; A.R.tab = dd->cut_table;
  if (!rule_lookup(dd, 1, order, &A.R.r1, &A.R.n1) || !rule_lookup(dd, 2, order, &A.R.r2, &A.R.n2)) {
    set_error("quad / hex rule: integration order beyond the tabulated reference rules");
    return NGSX_ERR_UNSUPPORTED;
  }
  A.err = reinterpret_cast<unsigned long long *>(err);
  if (err) { if (int rc = reset_err(err, st)) return rc; }
  return NGSX_OK;
}

}  // namespace ngsx

using namespace ngsx;

extern "C" {

int ngsx_q_classify(const ngsx_mesh *mesh, const double *lset_v, int8_t *dom, double *ratio, int64_t *err, void *stream) {
  QArgs A{};
  cudaStream_t st = as_stream(stream);
  if (int rc = q_args(mesh, lset_v, NEG, 0, err, st, A)) return rc;
  if (mesh->ne == 0) return NGSX_OK;
  const unsigned nb = (unsigned)((mesh->ne + 127) / 128);
  if (mesh->dim == 2) k_q_classify<2><<<nb, 128, 0, st>>>(A, dom, ratio);
  else k_q_classify<3><<<nb, 128, 0, st>>>(A, dom, ratio);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_q_rule_count(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, int32_t *npts, int64_t *err, void *stream) {
  QArgs A{};
  cudaStream_t st = as_stream(stream);
  if (int rc = q_args(mesh, lset_v, dt, order, err, st, A)) return rc;
  if (mesh->ne == 0) return NGSX_OK;
  NGSX_REQUIRE(npts, "q_rule_count: npts is NULL");
  const unsigned nb = (unsigned)((mesh->ne + 127) / 128);
  if (mesh->dim == 2) k_q_count<2><<<nb, 128, 0, st>>>(A, npts);
  else k_q_count<3><<<nb, 128, 0, st>>>(A, npts);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_q_rule_fill(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, int transform_if, const int64_t *offs,
                     double *pts3, double *wts, void *stream) {
  QArgs A{};
  cudaStream_t st = as_stream(stream);
  if (int rc = q_args(mesh, lset_v, dt, order, nullptr, st, A)) return rc;
  if (mesh->ne == 0) return NGSX_OK;
  NGSX_REQUIRE(offs && pts3 && wts, "q_rule_fill: NULL array");
  const unsigned nb = (unsigned)((mesh->ne + 127) / 128);
  if (mesh->dim == 2) k_q_fill<2><<<nb, 128, 0, st>>>(A, transform_if, offs, pts3, wts);
  else k_q_fill<3><<<nb, 128, 0, st>>>(A, transform_if, offs, pts3, wts);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_q_integrate(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, const ngsx_poly *h_f, double *out, int64_t *err,
                     void *stream) {
  QArgs A{};
  cudaStream_t st = as_stream(stream);
  if (int rc = q_args(mesh, lset_v, dt, order, err, st, A)) return rc;
  NGSX_REQUIRE(h_f && out && h_f->nterms >= 0 && h_f->nterms <= NGSX_POLY_MAXTERMS, "q_integrate: bad integrand");
  if (mesh->ne == 0) { NGSX_CUDA(cudaMemsetAsync(out, 0, sizeof(double), st)); return NGSX_OK; }
  QPoly P;
  P.n = h_f->nterms;
  for (int m = 0; m < P.n; m++) { P.c[m] = h_f->coef[m]; for (int d = 0; d < 3; d++) P.ex[m][d] = h_f->ex[m][d]; }
  const unsigned nb = (unsigned)((mesh->ne + 127) / 128);
  const int64_t nred = (int64_t)nb * 4;
  const unsigned nmid = (unsigned)std::min<int64_t>((nred + 255) / 256, 296);
  double *partial = nullptr;
  NGSX_CUDA(cudaMallocAsync(&partial, (nred + nmid) * sizeof(double), st));
  if (mesh->dim == 2) k_q_integrate<2><<<nb, 128, 0, st>>>(A, P, partial);
  else k_q_integrate<3><<<nb, 128, 0, st>>>(A, P, partial);
  cudaError_t e1 = cudaGetLastError();
  k_q_reduce<<<nmid, 256, 0, st>>>(nred, partial, partial + nred);
  k_q_final<<<1, 1024, 0, st>>>(nmid, partial + nred, out);
  g_launches.fetch_add(3);
  cudaError_t e2 = cudaGetLastError();
  cudaFreeAsync(partial, st);
  NGSX_CUDA(e1);
  NGSX_CUDA(e2);
  return NGSX_OK;
}

}  // extern "C"
