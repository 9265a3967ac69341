// ngsx_iso.cuh -- isoparametric (deformed) element geometry shared by ngsx_deformed.cu and ngsx_facetpatch.cu:
// x(xi) = sum_i c_i phi_i(xi) with the H1 basis of order DORD (vertex + edge functions), its Jacobian and inverse.
// Reference behaviour: ALE_ElementTransformation / FE_ElementTransformation::CalcJacobian (NGS/fem/elementtransformation.cpp).
#pragma once
#include "ngsx_device.cuh"

namespace ngsx {

template <int DIM>
__device__ __forceinline__ void invert(const double (*J)[DIM], double (*I)[DIM], double &det) {
  if constexpr (DIM == 2) {
    det = J[0][0] * J[1][1] - J[0][1] * J[1][0];
    double id = 1.0 / det;
    I[0][0] = id * J[1][1]; I[0][1] = -id * J[0][1];
    I[1][0] = -id * J[1][0]; I[1][1] = id * J[0][0];
  } else {
    double h0 = J[1][1] * J[2][2] - J[2][1] * J[1][2];
    double h1 = J[2][1] * J[0][2] - J[0][1] * J[2][2];
    double h2 = J[0][1] * J[1][2] - J[1][1] * J[0][2];
    det = J[0][0] * h0 + J[1][0] * h1 + J[2][0] * h2;
    double id = 1.0 / det;
    I[0][0] = id * h0; I[0][1] = id * h1; I[0][2] = id * h2;
    I[1][0] = id * (J[2][0] * J[1][2] - J[1][0] * J[2][2]);
    I[1][1] = id * (J[0][0] * J[2][2] - J[2][0] * J[0][2]);
    I[1][2] = id * (J[1][0] * J[0][2] - J[0][0] * J[1][2]);
    I[2][0] = id * (J[1][0] * J[2][1] - J[2][0] * J[1][1]);
    I[2][1] = id * (J[2][0] * J[0][1] - J[0][0] * J[2][1]);
    I[2][2] = id * (J[0][0] * J[1][1] - J[1][0] * J[0][1]);
  }
}

// geometry at a reference point: x, gl[k][:] = grad_x lam_k (rows of F^-1, last = minus their sum), det F
struct GcWarp {                       // coefficients of the warp's element, contiguous
  const double *p;
  __device__ __forceinline__ double operator()(int t) const { return p[t]; }
};
template <int STRIDE> struct GcThread {   // coefficients of the thread's element, thread-strided (conflict free)
  const double *p;
  __device__ __forceinline__ double operator()(int t) const { return p[t * STRIDE]; }
};

template <int DIM, int DORD, typename GC>
__device__ __forceinline__ void iso_point(const GC gc, const double *lam, double *x, double (*gl)[DIM], double &det) {
  constexpr int NDG = H1<DIM, DORD>::ND;
  double glr[DIM + 1][DIM];
#pragma unroll
  for (int k = 0; k <= DIM; k++)
#pragma unroll
    for (int d = 0; d < DIM; d++) glr[k][d] = (k == DIM) ? -1.0 : (k == d ? 1.0 : 0.0);
  double J[DIM][DIM];
#pragma unroll
  for (int r = 0; r < DIM; r++) {
    x[r] = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; k++) J[r][k] = 0.0;
  }
#pragma unroll
  for (int i = 0; i < NDG; i++) {
    double s, g[DIM];
    h1_shape_grad<DIM, DORD>(i, lam, glr, s, g);
#pragma unroll
    for (int r = 0; r < DIM; r++) {
      const double c = gc(i * DIM + r);
      x[r] += c * s;
#pragma unroll
      for (int k = 0; k < DIM; k++) J[r][k] += c * g[k];
    }
  }
  double I[DIM][DIM];
  invert<DIM>(J, I, det);
#pragma unroll
  for (int d = 0; d < DIM; d++) {
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; k++) { gl[k][d] = I[k][d]; s += I[k][d]; }
    gl[DIM][d] = -s;
  }
}

}  // namespace ngsx
