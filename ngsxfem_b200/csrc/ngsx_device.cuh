// ngsx_device.cuh -- device-side building blocks of the cut-FEM path (sm_100a).
// Geometry of affine simplices, the P1 level-set sign test, the sub-simplex decomposition of a cut simplex
// (table driven: one entry per vertex sign pattern), H1 high-order shape functions.
// Reference behaviour (what, not how): cutint/straightcutrule.cpp:11-204, 503-576; NGS/fem/h1hofe_impl.hpp:78-108, 250-312.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ngsx {

enum { NEG = 0, POS = 1, IF = 2 };

// ---------------------------------------------------------------------------------------------------------------
// Decomposition table.  A vertex of a sub-simplex is coded (i<<4)|j: i==j -> vertex i of the reference simplex,
// i<j -> the zero of the P1 level set on edge (i,j).  One entry per (dim, domain type, sign pattern); the sign
// pattern has bit k set when phi_k >= 0 (the reference's ">= 0 is POS" convention, straightcutrule.cpp:55,154).
// Built on the host by replaying the case analysis of LevelsetCutSimplex::Decompose (straightcutrule.cpp:131-196).
// ---------------------------------------------------------------------------------------------------------------
struct CutEntry {
  uint8_t ns;          // number of sub-simplices (0: pattern is not a cut)
  uint8_t npt;         // vertices per sub-simplex
  uint8_t v[3][4];
};
struct CutTable { CutEntry e[2][3][16]; };   // [dim-2][dt][mask]

template <int DIM> struct ElGeom {
  double Jinv[DIM][DIM];   // inverse Jacobian of x = sum_i lam_i p_i,  J[:,k] = p_k - p_DIM
  double det;              // det J (signed)
};

__device__ __forceinline__ double sel4(const double *a, int i) {
  return i == 0 ? a[0] : (i == 1 ? a[1] : (i == 2 ? a[2] : a[3]));
}

template <int DIM>
__device__ __forceinline__ void element_geometry(const double (*p)[DIM], ElGeom<DIM> &G) {
  if constexpr (DIM == 2) {
    double a = p[0][0] - p[2][0], b = p[1][0] - p[2][0];
    double c = p[0][1] - p[2][1], d = p[1][1] - p[2][1];
    double det = a * d - b * c;
    double id = 1.0 / det;
    G.det = det;
    G.Jinv[0][0] = id * d;  G.Jinv[0][1] = -id * b;
    G.Jinv[1][0] = -id * c; G.Jinv[1][1] = id * a;
  } else {
    double J[3][3];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
      for (int k = 0; k < 3; k++) J[r][k] = p[k][r] - p[3][r];
    double h0 = J[1][1] * J[2][2] - J[2][1] * J[1][2];
    double h1 = J[2][1] * J[0][2] - J[0][1] * J[2][2];
    double h2 = J[0][1] * J[1][2] - J[1][1] * J[0][2];
    double det = J[0][0] * h0 + J[1][0] * h1 + J[2][0] * h2;
    double id = 1.0 / det;
    G.det = det;
    G.Jinv[0][0] = id * h0; G.Jinv[0][1] = id * h1; G.Jinv[0][2] = id * h2;
    G.Jinv[1][0] = id * (J[2][0] * J[1][2] - J[1][0] * J[2][2]);
    G.Jinv[1][1] = id * (J[0][0] * J[2][2] - J[2][0] * J[0][2]);
    G.Jinv[1][2] = id * (J[1][0] * J[0][2] - J[0][0] * J[1][2]);
    G.Jinv[2][0] = id * (J[1][0] * J[2][1] - J[2][0] * J[1][1]);
    G.Jinv[2][1] = id * (J[2][0] * J[0][1] - J[0][0] * J[2][1]);
    G.Jinv[2][2] = id * (J[0][0] * J[1][1] - J[1][0] * J[0][1]);
  }
}

// sign test of CheckIfStraightCut (epsilon = 0).  returns: 0 NEG, 1 POS, 2 IF (cut), -1 degenerate
// (all zero, or strict mixed signs with an exact zero: the reference throws there).
// mask: bit k = (phi_k >= 0).
template <int DIM>
__device__ __forceinline__ int sign_test(const double *ls, int &mask) {
  bool haspos = false, hasneg = false, haszero = false;
  mask = 0;
#pragma unroll
  for (int k = 0; k <= DIM; k++) {
    double v = ls[k];
    haspos |= v > 0.0;
    hasneg |= v < 0.0;
    haszero |= v == 0.0;
    mask |= (v >= 0.0) ? (1 << k) : 0;
  }
  if (haspos && hasneg) return haszero ? -1 : IF;
  if (!haspos && !hasneg) return -1;
  return hasneg ? NEG : POS;
}

// reference-element coordinates of a coded sub-simplex vertex
template <int DIM>
__device__ __forceinline__ void coded_point(uint8_t code, const double *ls, double *P) {
  int i = code >> 4, j = code & 15;
  if (i == j) {
#pragma unroll
    for (int d = 0; d < 3; d++) P[d] = (d < DIM && d == i) ? 1.0 : 0.0;
  } else {
    double li = sel4(ls, i), lj = sel4(ls, j);
    double f = li / (li - lj);
#pragma unroll
    for (int d = 0; d < 3; d++) {
      double vi = (d < DIM && d == i) ? 1.0 : 0.0, vj = (d < DIM && d == j) ? 1.0 : 0.0;
      P[d] = vi + f * (vj - vi);
    }
  }
}

// measure factor of a sub-simplex with n points (SimpleX::GetVolume, straightcutrule.cpp:66-72)
__device__ __forceinline__ double simplex_factor(int n, const double (*P)[3]) {
  if (n == 1) return 1.0;
  if (n == 2) {
    double a0 = P[1][0] - P[0][0], a1 = P[1][1] - P[0][1], a2 = P[1][2] - P[0][2];
    return sqrt(a0 * a0 + a1 * a1 + a2 * a2);
  }
  if (n == 3) {
    double a0 = P[2][0] - P[0][0], a1 = P[2][1] - P[0][1], a2 = P[2][2] - P[0][2];
    double b0 = P[1][0] - P[0][0], b1 = P[1][1] - P[0][1], b2 = P[1][2] - P[0][2];
    double cx = a1 * b2 - a2 * b1, cy = a2 * b0 - a0 * b2, cz = a0 * b1 - a1 * b0;
    return sqrt(cx * cx + cy * cy + cz * cz);
  }
  double a0 = P[3][0] - P[0][0], a1 = P[3][1] - P[0][1], a2 = P[3][2] - P[0][2];
  double b0 = P[2][0] - P[0][0], b1 = P[2][1] - P[0][1], b2 = P[2][2] - P[0][2];
  double c0 = P[1][0] - P[0][0], c1 = P[1][1] - P[0][1], c2 = P[1][2] - P[0][2];
  double det = a0 * (b1 * c2 - b2 * c1) - b0 * (a1 * c2 - a2 * c1) + c0 * (a1 * b2 - a2 * b1);
  return fabs(det);
}

// point q of the plain rule on a sub-simplex (SimpleX::GetPlainIntegrationRule, straightcutrule.cpp:80-100):
// x = (1 - sum xi) P0 + sum_m xi_m P_{m+1}
__device__ __forceinline__ void simplex_point(int n, const double (*P)[3], const double4 r, double *x) {
  double xi[3] = {r.x, r.y, r.z};
  double ow = 1.0;
  for (int m = 0; m < n - 1; m++) ow -= xi[m];
#pragma unroll
  for (int d = 0; d < 3; d++) x[d] = ow * P[0][d];
  for (int m = 0; m < n - 1; m++)
#pragma unroll
    for (int d = 0; d < 3; d++) x[d] += xi[m] * P[m + 1][d];
}

// ---------------------------------------------------------------------------------------------------------------
// H1 high-order element, orders 1..2 (orientation free at these orders: edge function = -1/2 lam_a lam_b,
// IntLegNoBubble::P0, NGS/fem/recursive_pol.hpp:1372).  Local dof order: vertices, then edges in NGSolve's local
// edge order (NGS/fem/elementtopology.cpp:143-146 trig {2,0},{1,2},{0,1}; :161-167 tet {3,0},{3,1},{3,2},{0,1},{0,2},{1,2}).
// ---------------------------------------------------------------------------------------------------------------
template <int DIM, int ORDER> struct H1 {
  static constexpr int NEDGE = (DIM == 2) ? 3 : 6;
  static constexpr int NFACE = (DIM == 2) ? 1 : 4;        // order 3: one face / cell function lam_a lam_b lam_c each
  static constexpr int ND = (ORDER == 1) ? (DIM + 1) : (ORDER == 2 ? (DIM == 2 ? 6 : 10) : (DIM + 1 + 2 * NEDGE + NFACE));
};

// order 3 (element matrices only): per edge the pair {-1/2 lam_a lam_b, -1/2 x lam_a lam_b}, x = lam_e1 - lam_e0 with e0 the edge
// vertex of the SMALLER global number (vertex-oriented edge, NGS/fem/h1hofe_impl.hpp:78-108, 250-312; IntLegNoBubble P0 = -1/2,
// P1 = -x/2, NGS/fem/recursive_pol.hpp:1372-1374), then per face lam_a lam_b lam_c (DubinerBasis of order 0: orientation free;
// tet faces NGS/fem/elementtopology.cpp {3,1,2},{3,2,0},{3,0,1},{0,2,1}).  orient: bit e set = edge e runs against its local order.


template <int DIM> __device__ __forceinline__ void edge_vertices(int e, int &a, int &b) {
  if constexpr (DIM == 2) {
    a = (e == 0) ? 2 : (e == 1 ? 1 : 0);
    b = (e == 0) ? 0 : (e == 1 ? 2 : 1);
  } else {
    a = (e < 3) ? 3 : (e < 5 ? 0 : 1);
    b = (e < 3) ? e : (e == 3 ? 1 : 2);
  }
}

template <int DIM>
__device__ __forceinline__ unsigned edge_orientation(const int *v) {
  unsigned o = 0;
#pragma unroll
  for (int e = 0; e < H1<DIM, 3>::NEDGE; e++) {
    int a, b;
    edge_vertices<DIM>(e, a, b);
    const int va = a == 0 ? v[0] : (a == 1 ? v[1] : (a == 2 ? v[2] : v[DIM]));
    const int vb = b == 0 ? v[0] : (b == 1 ? v[1] : (b == 2 ? v[2] : v[DIM]));
    o |= (va > vb) ? (1u << e) : 0u;
  }
  return o;
}

// shape value of dof i at barycentrics lam
template <int DIM, int ORDER>
__device__ __forceinline__ double h1_shape(int i, const double *lam, unsigned orient = 0) {
  if (i <= DIM) return sel4(lam, i);
  constexpr int NEDGE = H1<DIM, ORDER>::NEDGE;
  constexpr int PER = ORDER == 3 ? 2 : 1;
  const int r = i - (DIM + 1);
  if (ORDER < 3 || r < PER * NEDGE) {
    const int e = r / PER, k = r % PER;
    int a, b;
    edge_vertices<DIM>(e, a, b);
    const double la = sel4(lam, a), lb = sel4(lam, b);
    if (ORDER == 3 && k == 1) return -0.5 * (((orient >> e) & 1u) ? la - lb : lb - la) * (la * lb);
    return -0.5 * la * lb;
  }
  const int f = r - PER * NEDGE;
  int a = 0, b = 1, c = 2;
  if (DIM == 3) { a = f == 3 ? 0 : 3; b = f == 0 ? 1 : (f == 1 ? 2 : (f == 2 ? 0 : 2)); c = f == 0 ? 2 : (f == 1 ? 0 : 1); }
  return sel4(lam, a) * sel4(lam, b) * sel4(lam, c);
}

// shape value + gradient w.r.t. the coordinates in which gl[k][:] (gradients of the barycentrics) are given
template <int DIM, int ORDER>
__device__ __forceinline__ void h1_shape_grad(int i, const double *lam, const double (*gl)[DIM], double &s, double *g, unsigned orient = 0) {
  if (i <= DIM) {
    s = sel4(lam, i);
#pragma unroll
    for (int d = 0; d < DIM; d++) g[d] = (i == 0 ? gl[0][d] : (i == 1 ? gl[1][d] : (i == 2 ? gl[2][d] : gl[DIM][d])));
    return;
  }
  constexpr int NEDGE = H1<DIM, ORDER>::NEDGE;
  constexpr int PER = ORDER == 3 ? 2 : 1;                 // functions per edge
  const int r = i - (DIM + 1);
  if (ORDER < 3 || r < PER * NEDGE) {
    const int e = r / PER, k = r % PER;
    int a, b;
    edge_vertices<DIM>(e, a, b);
    double la = sel4(lam, a), lb = sel4(lam, b);
    if (ORDER == 3 && k == 1) {
      const double sg = ((orient >> e) & 1u) ? -1.0 : 1.0;      // x = sg (lam_b - lam_a)
      const double c = la * lb, x = sg * (lb - la);
      s = -0.5 * x * c;
#pragma unroll
      for (int d = 0; d < DIM; d++) {
        double ga = (a == 0 ? gl[0][d] : (a == 1 ? gl[1][d] : (a == 2 ? gl[2][d] : gl[DIM][d])));
        double gb = (b == 0 ? gl[0][d] : (b == 1 ? gl[1][d] : (b == 2 ? gl[2][d] : gl[DIM][d])));
        g[d] = -0.5 * (sg * (gb - ga) * c + x * (ga * lb + la * gb));
      }
      return;
    }
    s = -0.5 * la * lb;
#pragma unroll
    for (int d = 0; d < DIM; d++) {
      double ga = (a == 0 ? gl[0][d] : (a == 1 ? gl[1][d] : (a == 2 ? gl[2][d] : gl[DIM][d])));
      double gb = (b == 0 ? gl[0][d] : (b == 1 ? gl[1][d] : (b == 2 ? gl[2][d] : gl[DIM][d])));
      g[d] = -0.5 * (ga * lb + la * gb);
    }
    return;
  }
  // order 3 face / cell function
  const int f = r - PER * NEDGE;
  int a = 0, b = 1, c = 2;
  if (DIM == 3) { a = f == 3 ? 0 : 3; b = f == 0 ? 1 : (f == 1 ? 2 : (f == 2 ? 0 : 2)); c = f == 0 ? 2 : (f == 1 ? 0 : 1); }
  const double la = sel4(lam, a), lb = sel4(lam, b), lc = sel4(lam, c);
  s = la * lb * lc;
#pragma unroll
  for (int d = 0; d < DIM; d++) {
    double ga = (a == 0 ? gl[0][d] : (a == 1 ? gl[1][d] : (a == 2 ? gl[2][d] : gl[DIM][d])));
    double gb = (b == 0 ? gl[0][d] : (b == 1 ? gl[1][d] : (b == 2 ? gl[2][d] : gl[DIM][d])));
    double gc = (c == 0 ? gl[0][d] : (c == 1 ? gl[1][d] : (c == 2 ? gl[2][d] : gl[DIM][d])));
    g[d] = ga * lb * lc + la * gb * lc + la * lb * gc;
  }
}

__device__ __forceinline__ double ipow(double x, int p) {
  double r = 1.0;
  if (p < 0) { x = 1.0 / x; p = -p; }
  for (int k = 0; k < p; k++) r *= x;
  return r;
}

}  // namespace ngsx
