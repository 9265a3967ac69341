/* TEST INFRASTRUCTURE -- CPU oracle, not part of the product path.
 *
 * Plain-C restatement of the reference algorithm for the cut-FEM hot path
 * (ngsxfem @ 88e7db38 + vendored NGSolve).  Every function cites the reference
 * file:line it follows (paths relative to /root/reference; NGS/ =
 * external_dependencies/ngsolve/).  Parity of this file is PINNED two ways
 * (tests/test_oracle_*.py):
 *   1. against the reference's own known-answer tests (tests/pytests/test_localgeom.py,
 *      test_straightcutrule.py, test_xfes_ndof.py), and
 *   2. against the reference's own sources compiled in place (oracle/_ref/libngsxref.so,
 *      built by oracle/build_ref.sh) and the committed golden vectors made from it
 *      (tests/golden/, tools/gen_golden.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
 * may load this library.  Scalar, single-threaded unless a *_mesh entry point says otherwise.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "rule_tables.h"

enum { NEG = 0, POS = 1, IF = 2 };          /* utils/ngsxstd.hpp:14-17 */
enum { ORC_NULL = -1, ORC_THROW = -2, ORC_CAP = -3, ORC_UNSUP = -4 };

/* ------------------------------------------------------------------ rule tables */
/* NGS/fem/intrule.cpp:2620-2659 SelectIntegrationRule (data dumped from the compiled reference) */
static int select_rule(int dim, int order, const double (**data)[4])
{
  if (order < 0) order = 0;
  switch (dim) {
  case 0: { static const double pt[1][4] = {{0, 0, 0, 1}}; *data = pt; return 1; }
  case 1: if (order > NGSX_RULE_SEGM_MAXORDER) return ORC_UNSUP;
    *data = NGSX_RULE_SEGM_DATA + NGSX_RULE_SEGM_OFFS[order]; return NGSX_RULE_SEGM_OFFS[order + 1] - NGSX_RULE_SEGM_OFFS[order];
  case 2: if (order > NGSX_RULE_TRIG_MAXORDER) return ORC_UNSUP;
    *data = NGSX_RULE_TRIG_DATA + NGSX_RULE_TRIG_OFFS[order]; return NGSX_RULE_TRIG_OFFS[order + 1] - NGSX_RULE_TRIG_OFFS[order];
  case 3: if (order > NGSX_RULE_TET_MAXORDER) return ORC_UNSUP;
    *data = NGSX_RULE_TET_DATA + NGSX_RULE_TET_OFFS[order]; return NGSX_RULE_TET_OFFS[order + 1] - NGSX_RULE_TET_OFFS[order];
  }
  return ORC_UNSUP;
}

int orc_select_rule(int dim, int order, double *pts3, double *wts, int cap)
{
  const double (*d)[4];
  int n = select_rule(dim, order, &d);
  if (n < 0) return n;
  if (n > cap) return ORC_CAP;
  for (int i = 0; i < n; i++) { pts3[3*i] = d[i][0]; pts3[3*i+1] = d[i][1]; pts3[3*i+2] = d[i][2]; wts[i] = d[i][3]; }
  return n;
}

/* ------------------------------------------------------------------ classification */
/* cutint/straightcutrule.cpp:11-39 CheckIfStraightCut (epsilon = 0) */
int orc_check_straight_cut(int n, const double *v)
{
  int haspos = 0, hasneg = 0;
  for (int i = 0; i < n; i++) {
    if (v[i] > 0.0) haspos = 1;
    if (v[i] < 0.0) hasneg = 1;
  }
  if ((hasneg && haspos) || (!hasneg && !haspos)) return IF;
  return hasneg ? NEG : POS;
}

/* ------------------------------------------------------------------ cut geometry */
typedef struct { double p[4][3]; int n; } simplex_t;       /* SimpleX: n points, D = n-1 */

static void ref_simplex(int dim, simplex_t *s)             /* cutint/straightcutrule.hpp:57-62 */
{
  memset(s, 0, sizeof *s);
  s->n = dim + 1;
  for (int i = 0; i < dim; i++) s->p[i][i] = 1.0;            /* (1,0,0),(0,1,0),(0,0,1),(0,0,0) */
}

/* cutint/straightcutrule.cpp:41-64 SimpleX::CalcIFPolytopEUsingLset.
 * returns number of cut points, or ORC_THROW where the reference throws
 * ("You tried to cut a simplex with a plain geometry lset function"). */
static int calc_cut_points(const simplex_t *s, const double *ls, double cut[6][3])
{
  if (orc_check_straight_cut(s->n, ls) != IF) return ORC_THROW;
  int D = s->n - 1;
  if (D == 1) {
    double f = ls[0] / (ls[0] - ls[1]);
    for (int d = 0; d < 3; d++) cut[0][d] = s->p[0][d] + f * (s->p[1][d] - s->p[0][d]);
    return 1;
  }
  int nc = 0;
  for (int i = 0; i < s->n; i++)
    for (int j = i + 1; j < s->n; j++)
      if ((ls[i] >= 0) != (ls[j] >= 0)) {
        double l2[2] = {ls[i], ls[j]};
        if (orc_check_straight_cut(2, l2) != IF) return ORC_THROW;     /* recursion into the 1-D SimpleX, :45-49 */
        double f = ls[i] / (ls[i] - ls[j]);
        for (int d = 0; d < 3; d++) cut[nc][d] = s->p[i][d] + f * (s->p[j][d] - s->p[i][d]);
        nc++;
      }
  return nc;
}

/* cutint/straightcutrule.cpp:66-72 SimpleX::GetVolume */
static double simplex_volume(const simplex_t *s)
{
  int D = s->n - 1;
  double a[3], b[3], c[3];
  if (D == 0) return 1.0;
  if (D == 1) {
    for (int d = 0; d < 3; d++) a[d] = s->p[1][d] - s->p[0][d];
    return sqrt(a[0]*a[0] + a[1]*a[1] + a[2]*a[2]);
  }
  if (D == 2) {
    for (int d = 0; d < 3; d++) { a[d] = s->p[2][d] - s->p[0][d]; b[d] = s->p[1][d] - s->p[0][d]; }
    double cx = a[1]*b[2] - a[2]*b[1], cy = a[2]*b[0] - a[0]*b[2], cz = a[0]*b[1] - a[1]*b[0];
    return sqrt(cx*cx + cy*cy + cz*cz);
  }
  for (int d = 0; d < 3; d++) { a[d] = s->p[3][d] - s->p[0][d]; b[d] = s->p[2][d] - s->p[0][d]; c[d] = s->p[1][d] - s->p[0][d]; }
  /* Determinant<3>(col0=a, col1=b, col2=c) */
  double det = a[0]*(b[1]*c[2] - b[2]*c[1]) - b[0]*(a[1]*c[2] - a[2]*c[1]) + c[0]*(a[1]*b[2] - a[2]*b[1]);
  return fabs(det);
}

/* cutint/straightcutrule.cpp:80-100 SimpleX::GetPlainIntegrationRule (appends) */
static int plain_rule(const simplex_t *s, int order, double *pts3, double *wts, int n0, int cap)
{
  const double (*r)[4];
  int D = s->n - 1;
  int nr = select_rule(D, order, &r);
  if (nr < 0) return nr;
  if (n0 + nr > cap) return ORC_CAP;
  double fac = simplex_volume(s);
  for (int q = 0; q < nr; q++) {
    double ow = 1.0;
    for (int m = 0; m < D; m++) ow -= r[q][m];
    double pt[3];
    for (int d = 0; d < 3; d++) pt[d] = ow * s->p[0][d];
    for (int m = 0; m < D; m++)
      for (int d = 0; d < 3; d++) pt[d] += r[q][m] * s->p[m + 1][d];
    for (int d = 0; d < 3; d++) pts3[3*(n0+q)+d] = pt[d];
    wts[n0 + q] = r[q][3] * fac;
  }
  return n0 + nr;
}

static void set_pts(simplex_t *s, int n, const double *a, const double *b, const double *c, const double *d)
{
  const double *q[4] = {a, b, c, d};
  s->n = n;
  for (int i = 0; i < n; i++) memcpy(s->p[i], q[i], 3 * sizeof(double));
}

/* cutint/straightcutrule.cpp:131-196 LevelsetCutSimplex::Decompose.
 * returns number of sub-simplices (<=3) or ORC_THROW. */
static int decompose(int dim, const double *ls, int dt, simplex_t out[3])
{
  simplex_t s;
  ref_simplex(dim, &s);
  double c[6][3];
  int nc = calc_cut_points(&s, ls, c);
  if (nc < 0) return nc;
  if (dt == IF) {
    if (nc == dim) {                                 /* s_cut is itself a (D-1)-simplex */
      out[0].n = nc;
      for (int i = 0; i < nc; i++) memcpy(out[0].p[i], c[i], sizeof c[i]);
      return 1;
    }
    if (nc == 4 && dim == 3) {
      set_pts(&out[0], 3, c[0], c[1], c[3], 0);
      set_pts(&out[1], 3, c[0], c[2], c[3], 0);
      return 2;
    }
    return ORC_THROW;                                 /* "Bad length of s_cut!" */
  }
  int rel[4], nrel = 0;
  for (int i = 0; i < dim + 1; i++)
    if ((dt == POS && ls[i] >= 0) || (dt == NEG && ls[i] < 0)) rel[nrel++] = i;
  if (nrel == 1) {                                    /* :156-160 {cuts..., v} */
    out[0].n = nc + 1;
    for (int i = 0; i < nc; i++) memcpy(out[0].p[i], c[i], sizeof c[i]);
    memcpy(out[0].p[nc], s.p[rel[0]], sizeof c[0]);
    if (out[0].n != dim + 1) return ORC_THROW;
    return 1;
  }
  if (nrel == 2 && dim == 2) {                        /* :161-168 */
    set_pts(&out[0], 3, s.p[rel[0]], s.p[rel[1]], c[1], 0);
    set_pts(&out[1], 3, c[0], c[1], s.p[rel[0]], 0);
    return 2;
  }
  if (nrel == 2 && dim == 3) {                        /* :169-178 */
    set_pts(&out[0], 4, s.p[rel[1]], c[1], c[2], c[3]);
    set_pts(&out[1], 4, s.p[rel[0]], s.p[rel[1]], c[1], c[2]);
    set_pts(&out[2], 4, c[0], c[1], c[2], s.p[rel[0]]);
    return 3;
  }
  if (nrel == 3 && dim == 3) {                        /* :179-188 */
    set_pts(&out[0], 4, c[0], c[1], c[2], s.p[rel[2]]);
    set_pts(&out[1], 4, s.p[rel[0]], s.p[rel[1]], s.p[rel[2]], c[1]);
    set_pts(&out[2], 4, c[0], c[1], s.p[rel[0]], s.p[rel[2]]);
    return 3;
  }
  return ORC_THROW;                                   /* "Cutting this part of a tetraeder is not implemented yet!" */
}

/* cutint/straightcutrule.cpp:663-726 StraightCutIntegrationRuleUntransformed (simplices).
 * returns npts >= 0, ORC_NULL (-1) for "no rule" (nullptr), ORC_THROW (-2) where the reference throws. */
int orc_cut_rule_untrafo(int dim, const double *ls, int dt, int order, double *pts3, double *wts, int cap)
{
  int dom = orc_check_straight_cut(dim + 1, ls);
  if (dom != IF) {
    if (dom != dt) return ORC_NULL;
    return orc_select_rule(dim, order, pts3, wts, cap);
  }
  simplex_t sub[3];
  int ns = decompose(dim, ls, dt, sub);
  if (ns < 0) return ns;
  int n = 0;
  for (int k = 0; k < ns; k++) {
    n = plain_rule(&sub[k], order, pts3, wts, n, cap);
    if (n < 0) return n;
  }
  return n;
}

/* ------------------------------------------------------------------ affine element geometry */
/* FE_ElementTransformation with the P1 element: x = sum_i lam_i p_i, lam = (xi..., 1-sum xi)
 * => J[:,k] = p_k - p_dim  (NGS/fem/elementtransformation.hpp:199-220; Appendix B of SURVEY.md) */
typedef struct { double J[3][3], Jinv[3][3], det, p0[3]; int dim; } affine_t;

static void affine_init(int dim, const double *pmat, affine_t *A)
{
  memset(A, 0, sizeof *A);
  A->dim = dim;
  for (int r = 0; r < dim; r++) {
    for (int k = 0; k < dim; k++) A->J[r][k] = pmat[k*dim + r] - pmat[dim*dim + r];
    A->p0[r] = pmat[dim*dim + r];
  }
  double (*J)[3] = A->J, (*I)[3] = A->Jinv;
  if (dim == 1) { A->det = J[0][0]; I[0][0] = 1.0 / J[0][0]; }
  else if (dim == 2) {
    double det = J[0][0]*J[1][1] - J[0][1]*J[1][0];
    A->det = det;
    double id = 1.0 / det;
    I[0][0] = id * J[1][1]; I[0][1] = -id * J[0][1]; I[1][0] = -id * J[1][0]; I[1][1] = id * J[0][0];
  } else {
    /* NGS/basiclinalg CalcInverse(Mat<3,3>): cofactors / det */
    double h0 = J[1][1]*J[2][2] - J[2][1]*J[1][2];
    double h1 = J[2][1]*J[0][2] - J[0][1]*J[2][2];
    double h2 = J[0][1]*J[1][2] - J[1][1]*J[0][2];
    double det = J[0][0]*h0 + J[1][0]*h1 + J[2][0]*h2;
    A->det = det;
    double id = 1.0 / det;
    I[0][0] = id*h0; I[0][1] = id*h1; I[0][2] = id*h2;
    I[1][0] = id*(J[2][0]*J[1][2] - J[1][0]*J[2][2]);
    I[1][1] = id*(J[0][0]*J[2][2] - J[2][0]*J[0][2]);
    I[1][2] = id*(J[1][0]*J[0][2] - J[0][0]*J[1][2]);
    I[2][0] = id*(J[1][0]*J[2][1] - J[2][0]*J[1][1]);
    I[2][1] = id*(J[2][0]*J[0][1] - J[0][0]*J[2][1]);
    I[2][2] = id*(J[0][0]*J[1][1] - J[1][0]*J[0][1]);
  }
}

static void affine_map(const affine_t *A, const double *xi, double *x)
{
  for (int r = 0; r < 3; r++) x[r] = 0;
  for (int r = 0; r < A->dim; r++) {
    double v = A->p0[r];
    for (int k = 0; k < A->dim; k++) v += A->J[r][k] * xi[k];
    x[r] = v;
  }
}

/* cutint/straightcutrule.cpp:503-537 LevelsetWrapper (simplex part): constant gradient of the P1 level set on the
 * reference element, normal = grad/|grad| */
static void lset_ref_normal(int dim, const double *ls, double *n)
{
  double g[3] = {0, 0, 0};
  for (int k = 0; k < dim; k++) g[k] = ls[k] - ls[dim];
  double nrm = sqrt(g[0]*g[0] + g[1]*g[1] + g[2]*g[2]);
  for (int k = 0; k < 3; k++) n[k] = g[k] / nrm;
}

/* cutint/straightcutrule.cpp:580-661 StraightCutIntegrationRule incl. :558-576
 * TransformQuadUntrafoToIRInterface: IF weights *= | F^-T n_ref |  (caller multiplies by |det F| afterwards) */
int orc_cut_rule(int dim, const double *pmat, const double *ls, int dt, int order, double *pts3, double *wts, int cap)
{
  int n = orc_cut_rule_untrafo(dim, ls, dt, order, pts3, wts, cap);
  if (n < 0) return n;
  if (dt == IF && orc_check_straight_cut(dim + 1, ls) == IF) {
    affine_t A;
    affine_init(dim, pmat, &A);
    double nr[3], v[3] = {0, 0, 0};
    lset_ref_normal(dim, ls, nr);
    for (int r = 0; r < dim; r++)
      for (int k = 0; k < dim; k++) v[r] += A.Jinv[k][r] * nr[k];     /* Trans(Finv) * n */
    double fac = sqrt(v[0]*v[0] + v[1]*v[1] + v[2]*v[2]);
    for (int i = 0; i < n; i++) wts[i] *= fac;
  }
  return n;
}

/* xfem/cutinfo.cpp:79-111: part volumes from two order-0 rules -> domain type + cut ratio */
int orc_cutinfo_element(int dim, const double *pmat, const double *ls, double *ratio)
{
  double pv[2] = {0, 0}, pts[3*8], w[8];
  const int order_np[2] = {POS, NEG};
  for (int k = 0; k < 2; k++) {
    int np = order_np[k];
    int n = orc_cut_rule(dim, pmat, ls, np, 0, pts, w, 8);
    if (n == ORC_THROW) return ORC_THROW;
    for (int i = 0; i < n; i++) pv[np] += w[i];
  }
  *ratio = pv[NEG] / (pv[NEG] + pv[POS]);
  if (pv[NEG] > 0.0) return pv[POS] > 0.0 ? IF : NEG;
  return POS;
}

/* ------------------------------------------------------------------ space-time cut rule */
#define ST_MAXTD 6
/* NodalTimeFE::CalcInterpolationPoints (spacetime/SpaceTimeFE.cpp:150-173) */
static void st_nodes(int ntd, double *nd)
{
  nd[0] = 0.0;
  switch (ntd - 1) {
    case 1: nd[1] = 1.0; break;
    case 2: nd[1] = 0.5; nd[2] = 1.0; break;
    case 3: nd[1] = 0.5*(1.0 - 1.0/sqrt(5.0)); nd[2] = 0.5*(1.0 + 1.0/sqrt(5.0)); nd[3] = 1.0; break;
    case 4: nd[1] = 0.5*(1.0 - sqrt(3.0/7.0)); nd[2] = 0.5; nd[3] = 0.5*(1.0 + sqrt(3.0/7.0)); nd[4] = 1.0; break;
    case 5: nd[1] = 0.5*(1.0 - sqrt(1.0/3.0 + 2.0*sqrt(7.0)/21.0)); nd[2] = 0.5*(1.0 - sqrt(1.0/3.0 - 2.0*sqrt(7.0)/21.0));
            nd[3] = 0.5*(1.0 + sqrt(1.0/3.0 - 2.0*sqrt(7.0)/21.0)); nd[4] = 0.5*(1.0 + sqrt(1.0/3.0 + 2.0*sqrt(7.0)/21.0));
            nd[5] = 1.0; break;
    default: break;
  }
}

/* NodalTimeFE::CalcShape / CalcDShape: Lagrange_Pol (spacetime/SpaceTimeFE.hpp:86-95) on a value + derivative pair (AutoDiff<1>) */
static void st_lagrange(int ntd, double x, double *sh, double *dsh)
{
  double nd[ST_MAXTD];
  st_nodes(ntd, nd);
  for (int i = 0; i < ntd; i++) {
    double val = 1.0, der = 0.0;
    for (int j = 0; j < ntd; j++)
      if (j != i) {
        double f = (x - nd[j]) / (nd[i] - nd[j]), df = 1.0 / (nd[i] - nd[j]);
        der = der * f + val * df;
        val *= f;
      }
    sh[i] = val;
    if (dsh) dsh[i] = der;
  }
}

static double st_eval(int ntd, const double *li, double xi)
{
  double sh[ST_MAXTD], v = 0;
  st_lagrange(ntd, xi, sh, NULL);
  for (int k = 0; k < ntd; k++) v += li[k] * sh[k];
  return v;
}

/* cutint/spacetimecutrule.cpp:6-96 root_finding: zeros in (0,1) of the level-set trace of ONE vertex in time; nodal time element
 * of order 0..5.  Orders 1, 2: the closed forms, same expressions as the reference, including the division by a = 0 when the
 * order-2 data is linear in time (IEEE inf / nan -> no root).  Orders >= 3 (:46-95): 50 subdivisions; a value that is exactly 0
 * at a subdivision point is a root; every sign change is refined by <= 70 rounds of "secant point with |value| < 5e-16 ends,
 * else bisection step"; roots in the order "exact zeros first, then the refined intervals" (values are sorted by the caller). */
static int st_roots(int ntd, const double *li, double *roots)
{
  int n = 0;
  if (ntd == 2) {
    if ((li[0] >= 0) != (li[1] >= 0)) roots[n++] = -li[0] / (li[1] - li[0]);
  } else if (ntd == 3) {
    double c = li[0], a = 2*li[0] + 2*li[2] - 4.*li[1], b = li[2] - a - c;
    const double w = b*b - 4*a*c;
    if (fabs(w) < 1e-12) {
      double x = -b / (2*a);
      if (x < 1 && x > 0) roots[n++] = x;
    } else {
      double x1 = (b + sqrt(w)) / (-2.0*a), x2 = (b - sqrt(w)) / (-2.0*a);
      if (x1 < 1 && x1 > 0) roots[n++] = x1;
      if (x2 < 1 && x2 > 0) roots[n++] = x2;
    }
  } else if (ntd > 3) {
    const int subdivs = 50, bisection_iterations = 70;
    const double delta_x = 1./subdivs;
    double vals[51], ia[51];
    int ni = 0;
    for (int i = 0; i < subdivs + 1; i++) {
      double xi = delta_x * i;
      vals[i] = st_eval(ntd, li, xi);
      if (vals[i] == 0 && n < ST_MAXTD - 1) roots[n++] = xi;
      if (i >= 1) if (vals[i-1] * vals[i] < 0) ia[ni++] = xi - delta_x;
    }
    for (int m = 0; m < ni && n < ST_MAXTD - 1; m++) {
      double a = ia[m], b = ia[m] + delta_x, x_mid;
      double aval = st_eval(ntd, li, a), bval = st_eval(ntd, li, b);
      for (int j = 0; j < bisection_iterations; j++) {
        const double x_lin = a - aval*(b - a)/(bval - aval);
        double val = st_eval(ntd, li, x_lin);
        if (2*fabs(val) < 1e-15) { a = b = x_lin; break; }
        x_mid = 0.5*(a + b);
        val = st_eval(ntd, li, x_mid);
        if (val == 0) break;
        if (val * aval < 0) { b = x_mid; bval = val; }
        else if (val * bval < 0) { a = x_mid; aval = val; }
        else break;                                    /* reference: throws "Strange sign structure during bisection!" */
      }
      roots[n++] = 0.5*(a + b);
    }
  }
  return n;
}

static void st_time_shape(int ntd, double t, double *sh)
{
  if (ntd == 1) sh[0] = 1.0;
  else if (ntd == 2) { sh[0] = 1.0 - t; sh[1] = t; }
  else if (ntd == 3) { sh[0] = (1 - t) * (1 - 2*t) ; sh[1] = 4*t*(1 - t); sh[2] = t*(2*t - 1); }
  else st_lagrange(ntd, t, sh, NULL);
}

/* cutint/spacetimecutrule.cpp:98-196 SpaceTimeCutIntegrationRule.  lset_st [ntd][dim+1] (time-major, SpaceTimeFESpace::GetDofNrs);
 * out: spatial reference points, reference time per point, weights (wei_arr: w_space * w_time * (t1 - t0)).
 * returns the number of points, ORC_NULL for an empty rule. */
int orc_st_cut_rule(int dim, const double *pmat, const double *lset_st, int ntd, int dt, int order_time, int order_space,
                    double *pts3, double *tq, double *wts, int cap)
{
  if (ntd < 1 || ntd > ST_MAXTD) return ORC_UNSUP;
  const int nv = dim + 1;
  double cp[2 + 4*ST_MAXTD];
  int ncp = 0;
  cp[ncp++] = 0.0; cp[ncp++] = 1.0;
  for (int i = 0; i < nv; i++) {
    double li[ST_MAXTD], r[ST_MAXTD];
    for (int k = 0; k < ntd; k++) li[k] = lset_st[k*nv + i];
    int nr = st_roots(ntd, li, r);
    for (int k = 0; k < nr; k++) cp[ncp++] = r[k];
  }
  for (int i = 1; i < ncp; i++) {                      /* sort (:121) */
    double v = cp[i]; int j = i - 1;
    while (j >= 0 && cp[j] > v) { cp[j+1] = cp[j]; j--; }
    cp[j+1] = v;
  }
  const double (*rt)[4];
  int nrt = select_rule(1, order_time, &rt);
  if (nrt < 0) return nrt;
  int n = 0;
  for (int i = 0; i + 1 < ncp; i++) {
    const double t0 = cp[i], t1 = cp[i+1];
    for (int q = 0; q < nrt; q++) {
      const double t = t0 + rt[q][0] * (t1 - t0);
      double sh[ST_MAXTD], ls[4];
      st_time_shape(ntd, t, sh);
      for (int v = 0; v < nv; v++) {
        double d = 0;
        for (int k = 0; k < ntd; k++) d += lset_st[k*nv + v] * sh[k];
        if (fabs(d) < 1e-14) d = 1e-14;               /* :135 */
        ls[v] = d;
      }
      int dom = orc_check_straight_cut(nv, ls);
      int m = 0;
      if (dom == IF) m = orc_cut_rule(dim, pmat, ls, dt, order_space, pts3 + 3*n, wts + n, cap - n);
      else if (dom == dt) m = orc_select_rule(dim, order_space, pts3 + 3*n, wts + n, cap - n);
      if (m == ORC_CAP || m == ORC_THROW || m == ORC_UNSUP) return m;
      if (m < 0) m = 0;
      for (int k = n; k < n + m; k++) { wts[k] = wts[k] * rt[q][3] * (t1 - t0); tq[k] = t; }
      n += m;
    }
  }
  return n == 0 ? ORC_NULL : n;
}

/* ------------------------------------------------------------------ H1 high-order shapes, orders 1..3 */
/* NGS/fem/elementtopology.cpp:143-146 (trig edges), :161-167 (tet edges), faces :171-176 */
static const int TRIG_EDGES[3][2] = {{2,0},{1,2},{0,1}};
static const int TET_EDGES[6][2] = {{3,0},{3,1},{3,2},{0,1},{0,2},{1,2}};
static const int TET_FACES[4][3] = {{3,1,2},{3,2,0},{3,0,1},{0,2,1}};

int orc_h1_ndof(int dim, int order)
{
  if (dim == 2) return (order + 1) * (order + 2) / 2;
  if (dim == 3) return (order + 1) * (order + 2) * (order + 3) / 6;
  return order + 1;
}

/* NGS/fem/h1hofe_impl.hpp:78-108 (trig), :250-312 (tet); IntLegNoBubble P0=-1/2, P1=-x/2 (NGS/fem/recursive_pol.hpp:1372-1374);
 * vertex-oriented edge: e[0],e[1] swapped so that vnums[e[0]] > vnums[e[1]]  (NGS/fem/elementtopology.hpp GetVertexOrientedEdge).
 * shape[ndof], dshape[ndof][dim] w.r.t. reference coordinates. */
int orc_h1_shape(int dim, int order, const int *vnums, const double *xi, double *shape, double *dshape)
{
  if (order < 1 || order > 3 || dim < 2 || dim > 3) return ORC_UNSUP;
  int nv = dim + 1;
  double lam[4], dlam[4][3];
  memset(dlam, 0, sizeof dlam);
  double s = 0;
  for (int i = 0; i < dim; i++) { lam[i] = xi[i]; dlam[i][i] = 1.0; s += xi[i]; }
  lam[dim] = 1.0 - s;
  for (int d = 0; d < dim; d++) dlam[dim][d] = -1.0;
  int ii = 0;
  for (int i = 0; i < nv; i++, ii++) {
    shape[ii] = lam[i];
    for (int d = 0; d < dim; d++) dshape[ii*dim + d] = dlam[i][d];
  }
  if (order >= 2) {
    int ne = dim == 2 ? 3 : 6;
    for (int e = 0; e < ne; e++) {
      int a = dim == 2 ? TRIG_EDGES[e][0] : TET_EDGES[e][0];
      int b = dim == 2 ? TRIG_EDGES[e][1] : TET_EDGES[e][1];
      int va = vnums ? vnums[a] : a, vb = vnums ? vnums[b] : b;
      if (va > vb) { int t = a; a = b; b = t; }          /* e[0] = vertex with the smaller global number */
      /* EvalScaledMult(n, x = lam[e1]-lam[e0], t = lam[e0]+lam[e1], c = lam[e0]*lam[e1]):  c*P0, c*P1(x) , ... */
      double c = lam[a] * lam[b];
      double x = lam[b] - lam[a];
      shape[ii] = -0.5 * c;
      for (int d = 0; d < dim; d++) dshape[ii*dim + d] = -0.5 * (dlam[a][d]*lam[b] + lam[a]*dlam[b][d]);
      ii++;
      if (order >= 3) {
        shape[ii] = -0.5 * x * c;
        for (int d = 0; d < dim; d++)
          dshape[ii*dim + d] = -0.5 * ((dlam[b][d] - dlam[a][d]) * c + x * (dlam[a][d]*lam[b] + lam[a]*dlam[b][d]));
        ii++;
      }
    }
  }
  if (order >= 3) {
    int nf = dim == 2 ? 1 : 4;
    for (int f = 0; f < nf; f++) {
      int a = dim == 2 ? 0 : TET_FACES[f][0], b = dim == 2 ? 1 : TET_FACES[f][1], c = dim == 2 ? 2 : TET_FACES[f][2];
      /* DubinerBasis order 0 = 1  times  lam_a lam_b lam_c */
      shape[ii] = lam[a] * lam[b] * lam[c];
      for (int d = 0; d < dim; d++)
        dshape[ii*dim + d] = dlam[a][d]*lam[b]*lam[c] + lam[a]*dlam[b][d]*lam[c] + lam[a]*lam[b]*dlam[c][d];
      ii++;
    }
  }
  return ii;
}

/* ------------------------------------------------------------------ element matrix / vector */
#define ORC_MAXPTS 1024
#define ORC_MAXND 20

static double eval_mono(int nmono, const double *mono, const double *x)
{
  double f = 0;
  for (int m = 0; m < nmono; m++)
    f += mono[4*m] * pow(x[0], mono[4*m+1]) * pow(x[1], mono[4*m+2]) * pow(x[2], mono[4*m+3]);
  return f;
}

/* xfem/symboliccutbfi.cpp:73-315 T_CalcElementMatrixAdd for  form = c_stiff grad(u).grad(v) + c_mass u v  on an affine simplex:
 * rule :117, D = coef*measure*w :219-227, elmat += B D B^T :246-301.  elmat is ndof x ndof row-major (rows = test). */
int orc_elmat_h1(int dim, int fe_order, const int *vnums, const double *pmat, const double *ls, int dt, int intorder,
                 double c_mass, double c_stiff, double *elmat)
{
  static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS];
  int np = orc_cut_rule(dim, pmat, ls, dt, intorder, pts, w, ORC_MAXPTS);
  int nd = orc_h1_ndof(dim, fe_order);
  if (np < 0) return np;
  affine_t A;
  affine_init(dim, pmat, &A);
  double meas = fabs(A.det);
  for (int i = 0; i < nd*nd; i++) elmat[i] = 0;
  double shape[ORC_MAXND], dref[ORC_MAXND*3], dphys[ORC_MAXND*3];
  for (int q = 0; q < np; q++) {
    if (orc_h1_shape(dim, fe_order, vnums, pts + 3*q, shape, dref) < 0) return ORC_UNSUP;
    for (int i = 0; i < nd; i++)                     /* dshape_phys = dshape_ref * Finv  (DiffOpGradient) */
      for (int d = 0; d < dim; d++) {
        double v = 0;
        for (int k = 0; k < dim; k++) v += dref[i*dim + k] * A.Jinv[k][d];
        dphys[i*dim + d] = v;
      }
    double fac = meas * w[q];
    for (int i = 0; i < nd; i++)
      for (int j = 0; j < nd; j++) {
        double g = 0;
        for (int d = 0; d < dim; d++) g += dphys[i*dim + d] * dphys[j*dim + d];
        elmat[i*nd + j] += fac * (c_stiff * g + c_mass * shape[i] * shape[j]);
      }
  }
  return np;
}

/* T_CalcElementMatrixAdd with a GridFunction coefficient multiplying the form: the reference evaluates the CoefficientFunction tree per point
 * (xfem/symboliccutbfi.cpp:178-215); a GridFunction CF is  sum_i cf_local[i] phi_i  with the H1 shapes of its own order. */
int orc_elmat_h1_cf(int dim, int fe_order, const int *vnums, const double *pmat, const double *ls, int dt, int intorder,
                    double c_mass, double c_stiff, int cf_order, const double *cf_local, double *elmat)
{
  static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS];
  int np = orc_cut_rule(dim, pmat, ls, dt, intorder, pts, w, ORC_MAXPTS);
  int nd = orc_h1_ndof(dim, fe_order), ncf = orc_h1_ndof(dim, cf_order);
  for (int i = 0; i < nd*nd; i++) elmat[i] = 0;
  if (np < 0) return np;
  affine_t A;
  affine_init(dim, pmat, &A);
  double meas = fabs(A.det);
  double shape[ORC_MAXND], cshape[ORC_MAXND], dref[ORC_MAXND*3], cdref[ORC_MAXND*3], dphys[ORC_MAXND*3];
  for (int q = 0; q < np; q++) {
    if (orc_h1_shape(dim, fe_order, vnums, pts + 3*q, shape, dref) < 0) return ORC_UNSUP;
    if (orc_h1_shape(dim, cf_order, vnums, pts + 3*q, cshape, cdref) < 0) return ORC_UNSUP;
    double g = 0;
    for (int i = 0; i < ncf; i++) g += cf_local[i] * cshape[i];
    for (int i = 0; i < nd; i++)
      for (int d = 0; d < dim; d++) {
        double v = 0;
        for (int k = 0; k < dim; k++) v += dref[i*dim + k] * A.Jinv[k][d];
        dphys[i*dim + d] = v;
      }
    double fac = meas * w[q] * g;
    for (int i = 0; i < nd; i++)
      for (int j = 0; j < nd; j++) {
        double gg = 0;
        for (int d = 0; d < dim; d++) gg += dphys[i*dim + d] * dphys[j*dim + d];
        elmat[i*nd + j] += fac * (c_stiff * gg + c_mass * shape[i] * shape[j]);
      }
  }
  return np;
}

/* xfem/symboliccutlfi.cpp:52-158 T_CalcElementVector, f = sum_m c_m x^a y^b z^c */
int orc_elvec_h1(int dim, int fe_order, const int *vnums, const double *pmat, const double *ls, int dt, int intorder,
                 int nmono, const double *mono, double *elvec)
{
  static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS];
  int np = orc_cut_rule(dim, pmat, ls, dt, intorder, pts, w, ORC_MAXPTS);
  int nd = orc_h1_ndof(dim, fe_order);
  if (np < 0) return np;
  affine_t A;
  affine_init(dim, pmat, &A);
  double meas = fabs(A.det);
  for (int i = 0; i < nd; i++) elvec[i] = 0;
  double shape[ORC_MAXND], dref[ORC_MAXND*3], x[3];
  for (int q = 0; q < np; q++) {
    orc_h1_shape(dim, fe_order, vnums, pts + 3*q, shape, dref);
    affine_map(&A, pts + 3*q, x);
    double pv = meas * w[q] * eval_mono(nmono, mono, x);
    for (int i = 0; i < nd; i++) elvec[i] += pv * shape[i];
  }
  return np;
}

/* the same with a GridFunction factor in the coefficient: f(x) = poly(x) * gf(x).  The reference evaluates the CoefficientFunction tree
 * per point (xfem/symboliccutlfi.cpp:120-140); a GridFunction CF on the element is  sum_i cf_local[i] phi_i  with the H1 shapes of its
 * own order at the same reference point (NGS/comp/gridfunction.cpp GridFunctionCoefficientFunction::Evaluate). */
int orc_elvec_h1_cf(int dim, int fe_order, const int *vnums, const double *pmat, const double *ls, int dt, int intorder,
                    int nmono, const double *mono, int cf_order, const double *cf_local, double *elvec)
{
  static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS];
  int np = orc_cut_rule(dim, pmat, ls, dt, intorder, pts, w, ORC_MAXPTS);
  int nd = orc_h1_ndof(dim, fe_order), ncf = orc_h1_ndof(dim, cf_order);
  for (int i = 0; i < nd; i++) elvec[i] = 0;
  if (np < 0) return np;
  affine_t A;
  affine_init(dim, pmat, &A);
  double meas = fabs(A.det);
  double shape[ORC_MAXND], cshape[ORC_MAXND], dref[ORC_MAXND*3], x[3];
  for (int q = 0; q < np; q++) {
    orc_h1_shape(dim, fe_order, vnums, pts + 3*q, shape, dref);
    if (orc_h1_shape(dim, cf_order, vnums, pts + 3*q, cshape, dref) < 0) return ORC_UNSUP;
    affine_map(&A, pts + 3*q, x);
    double g = 0;
    for (int i = 0; i < ncf; i++) g += cf_local[i] * cshape[i];
    double pv = meas * w[q] * eval_mono(nmono, mono, x) * g;
    for (int i = 0; i < nd; i++) elvec[i] += pv * shape[i];
  }
  return np;
}

/* ------------------------------------------------------------------ deformed (isoparametric) elements, SURVEY 8 f-4 */
/* mesh.SetDeformation(gf) (NGS/comp/meshaccess.cpp ALE_ElementTransformation: point and Jacobian of the base
 * transformation plus the deformation field and its reference gradient) == FE_ElementTransformation with the H1 element of
 * the deformation's order and pointmat = coefficients of x(xi) in that basis (NGS/fem/elementtransformation.cpp
 * CalcJacobian: dxdxi = pointmat * dshape; CalcPoint: pointmat * shape):
 *   gcoef[ndg][dim]: vertex dofs = vertex coordinates + deformation, edge dofs (order 2) = deformation.
 * out: x, J[r][k] = dx_r/dxi_k, Jinv, det */
static void iso_eval(int dim, int dord, const double *gcoef, const double *xi, double *x, double J[3][3], double Jinv[3][3], double *det)
{
  double sh[ORC_MAXND], dsh[ORC_MAXND*3];
  int ndg = orc_h1_shape(dim, dord, 0, xi, sh, dsh);
  double pm[12];
  affine_t A;
  memset(J, 0, 9*sizeof(double));
  for (int r = 0; r < 3; r++) x[r] = 0;
  for (int i = 0; i < ndg; i++)
    for (int r = 0; r < dim; r++) {
      x[r] += gcoef[i*dim + r] * sh[i];
      for (int k = 0; k < dim; k++) J[r][k] += gcoef[i*dim + r] * dsh[i*dim + k];
    }
  /* reuse the affine inverse (NGS/basiclinalg CalcInverse) on a simplex whose edge vectors are the columns of J */
  for (int k = 0; k < dim; k++) for (int r = 0; r < dim; r++) pm[k*dim + r] = J[r][k];
  for (int r = 0; r < dim; r++) pm[dim*dim + r] = 0.0;
  affine_init(dim, pm, &A);
  memcpy(Jinv, A.Jinv, sizeof A.Jinv);
  *det = A.det;
}

/* cutint/straightcutrule.cpp:580-661 with a non-affine trafo: points/weights of the untransformed rule; IF weights times
 * |F(xi_q)^-T n_ref| per point (:558-576).  Also returns per point the measure |det F| and the mapped point. */
int orc_cut_rule_iso(int dim, int dord, const double *gcoef, const double *ls, int dt, int order,
                     double *pts3, double *wts, double *meas, double *xphys, int cap)
{
  int n = orc_cut_rule_untrafo(dim, ls, dt, order, pts3, wts, cap);
  if (n < 0) return n;
  int iface = dt == IF && orc_check_straight_cut(dim + 1, ls) == IF;
  double nr[3];
  if (iface) lset_ref_normal(dim, ls, nr);
  for (int q = 0; q < n; q++) {
    double x[3], J[3][3], Ji[3][3], det;
    iso_eval(dim, dord, gcoef, pts3 + 3*q, x, J, Ji, &det);
    if (iface) {
      double v[3] = {0, 0, 0};
      for (int r = 0; r < dim; r++)
        for (int k = 0; k < dim; k++) v[r] += Ji[k][r] * nr[k];
      wts[q] *= sqrt(v[0]*v[0] + v[1]*v[1] + v[2]*v[2]);
    }
    if (meas) meas[q] = fabs(det);
    if (xphys) for (int r = 0; r < 3; r++) xphys[3*q + r] = x[r];
  }
  return n;
}

/* xfem/symboliccutbfi.cpp:73-315 on a deformed element.  form[8] = {c_mass, c_stiff, c_nitsche, c_dndn, hpow_mass, hpow_stiff,
 * hpow_nitsche, hpow_dndn}: c_m u v + c_s grad u.grad v + c_n ((grad u.n) v + (grad v.n) u) + c_d (grad u.n)(grad v.n),
 * n = grad(lset)/|grad(lset)| evaluated per point on the deformed element (a GridFunction CF), h = specialcf.mesh_size =
 * |det F|^(1/dim) (NGS/fem/python_fem.cpp:564-593) evaluated PER POINT: MeshSizeCF is not flagged ElementwiseConstant
 * (NGS/fem/coefficient.hpp:22 default false; products inherit c1 && c2, NGS/fem/coefficient.cpp:1349), so the reference takes
 * the cf->Evaluate(mir, val) branch of :205-214.  Pinned against ref_elmat_h1_form (oracle/ref_driver.cpp). */
int orc_elmat_h1_iso(int dim, int fe_order, int dord, const double *gcoef, const double *ls, int dt, int intorder,
                     const double *form, double *elmat)
{
  static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS], meas[ORC_MAXPTS];
  int np = orc_cut_rule_iso(dim, dord, gcoef, ls, dt, intorder, pts, w, meas, 0, ORC_MAXPTS);
  int nd = orc_h1_ndof(dim, fe_order);
  if (np < 0) return np;
  for (int i = 0; i < nd*nd; i++) elmat[i] = 0;
  double shape[ORC_MAXND], dref[ORC_MAXND*3], dphys[ORC_MAXND*3], dn[ORC_MAXND];
  for (int q = 0; q < np; q++) {
    double c[4] = {form[0], form[1], form[2], form[3]};
    {
      double h = pow(meas[q], 1.0 / dim);
      for (int t = 0; t < 4; t++) {
        int p = (int)form[4 + t];
        double f = 1.0, b = p < 0 ? 1.0 / h : h;
        for (int k = 0; k < abs(p); k++) f *= b;
        c[t] *= f;
      }
    }
    double x[3], J[3][3], Ji[3][3], det, n[3] = {0, 0, 0};
    iso_eval(dim, dord, gcoef, pts + 3*q, x, J, Ji, &det);
    if (orc_h1_shape(dim, fe_order, 0, pts + 3*q, shape, dref) < 0) return ORC_UNSUP;
    for (int i = 0; i < nd; i++)
      for (int d = 0; d < dim; d++) {
        double v = 0;
        for (int k = 0; k < dim; k++) v += dref[i*dim + k] * Ji[k][d];
        dphys[i*dim + d] = v;
      }
    double g2 = 0;
    for (int d = 0; d < dim; d++) {
      for (int k = 0; k < dim; k++) n[d] += (ls[k] - ls[dim]) * Ji[k][d];      /* grad of the P1 level set */
      g2 += n[d]*n[d];
    }
    for (int d = 0; d < dim; d++) n[d] /= sqrt(g2);
    for (int i = 0; i < nd; i++) {
      dn[i] = 0;
      for (int d = 0; d < dim; d++) dn[i] += dphys[i*dim + d] * n[d];
    }
    double fac = meas[q] * w[q];
    for (int i = 0; i < nd; i++)
      for (int j = 0; j < nd; j++) {
        double g = 0;
        for (int d = 0; d < dim; d++) g += dphys[i*dim + d] * dphys[j*dim + d];
        elmat[i*nd + j] += fac * (c[1] * g + c[0] * shape[i] * shape[j] + c[2] * (dn[j]*shape[i] + dn[i]*shape[j]) + c[3] * dn[i]*dn[j]);
      }
  }
  return np;
}

/* xfem/symboliccutlfi.cpp:52-158 T_CalcElementVector on a deformed element, f = sum_m c_m x^a y^b z^c at the mapped point */
int orc_elvec_h1_iso(int dim, int fe_order, int dord, const double *gcoef, const double *ls, int dt, int intorder,
                     int nmono, const double *mono, double *elvec)
{
  static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS], meas[ORC_MAXPTS], xp[3*ORC_MAXPTS];
  int np = orc_cut_rule_iso(dim, dord, gcoef, ls, dt, intorder, pts, w, meas, xp, ORC_MAXPTS);
  int nd = orc_h1_ndof(dim, fe_order);
  if (np < 0) return np;
  for (int i = 0; i < nd; i++) elvec[i] = 0;
  double shape[ORC_MAXND], dref[ORC_MAXND*3];
  for (int q = 0; q < np; q++) {
    orc_h1_shape(dim, fe_order, 0, pts + 3*q, shape, dref);
    double pv = meas[q] * w[q] * eval_mono(nmono, mono, xp + 3*q);
    for (int i = 0; i < nd; i++) elvec[i] += pv * shape[i];
  }
  return np;
}

/* coefficients of x(xi) on element e: vertex coordinates + deformation on the vertex dofs, deformation on the others */
static void gather_iso(int dim, const int32_t *el2v, const double *vert, const double *lset_v, int ndg, const int32_t *el2dof_def,
                       const double *defo, int64_t e, double *gcoef, double *ls)
{
  int nv = dim + 1;
  for (int i = 0; i < ndg; i++) {
    int64_t g = el2dof_def[e*ndg + i];
    for (int d = 0; d < dim; d++) gcoef[i*dim + d] = defo[g*dim + d];
  }
  for (int i = 0; i < nv; i++) {
    int64_t v = el2v[e*nv + i];
    ls[i] = lset_v[v];
    for (int d = 0; d < dim; d++) gcoef[i*dim + d] += vert[v*dim + d];
  }
}

/* IntegrateX (cutint/python_cutint.cpp:20-72) on a deformed mesh, cf = polynomial in the mapped point */
double orc_integrate_mesh_iso(int dim, int64_t ne, const int32_t *el2v, const double *vert, const double *lset_v,
                              int dord, const int32_t *el2dof_def, const double *defo, int dt, int order,
                              int nmono, const double *mono, int64_t *npts_total)
{
  double sum = 0;
  int64_t nptot = 0;
  int ndg = orc_h1_ndof(dim, dord);
#pragma omp parallel for schedule(static) reduction(+:sum,nptot)
  for (int64_t e = 0; e < ne; e++) {
    static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS], meas[ORC_MAXPTS], xp[3*ORC_MAXPTS];
    double gcoef[ORC_MAXND*3], ls[4];
    gather_iso(dim, el2v, vert, lset_v, ndg, el2dof_def, defo, e, gcoef, ls);
    int np = orc_cut_rule_iso(dim, dord, gcoef, ls, dt, order, pts, w, meas, xp, ORC_MAXPTS);
    if (np < 0) continue;
    double lsum = 0;
    for (int q = 0; q < np; q++) lsum += meas[q] * w[q] * eval_mono(nmono, mono, xp + 3*q);
    sum += lsum;
    nptot += np;
  }
  if (npts_total) *npts_total = nptot;
  return sum;
}

int64_t orc_elvec_mesh_iso(int dim, int64_t nel, const int64_t *els, const int32_t *el2v, const double *vert, const double *lset_v,
                           int dord, const int32_t *el2dof_def, const double *defo, int fe_order, int dt, int intorder,
                           int nmono, const double *mono, double *elvecs, int32_t *npts)
{
  int nd = orc_h1_ndof(dim, fe_order), ndg = orc_h1_ndof(dim, dord);
  int64_t nn = 0;
#pragma omp parallel for schedule(dynamic,256) reduction(+:nn)
  for (int64_t k = 0; k < nel; k++) {
    int64_t e = els ? els[k] : k;
    double gcoef[ORC_MAXND*3], ls[4];
    gather_iso(dim, el2v, vert, lset_v, ndg, el2dof_def, defo, e, gcoef, ls);
    int r = orc_elvec_h1_iso(dim, fe_order, dord, gcoef, ls, dt, intorder, nmono, mono, elvecs + (size_t)k*nd);
    if (npts) npts[k] = r;
    if (r < 0) memset(elvecs + (size_t)k*nd, 0, sizeof(double)*nd); else nn++;
  }
  return nn;
}

int64_t orc_elmat_mesh_iso(int dim, int64_t nel, const int64_t *els, const int32_t *el2v, const double *vert, const double *lset_v,
                           int dord, const int32_t *el2dof_def, const double *defo, int fe_order, int dt, int intorder,
                           const double *form, double *elmats, int32_t *npts)
{
  int nd = orc_h1_ndof(dim, fe_order), ndg = orc_h1_ndof(dim, dord);
  int64_t nn = 0;
#pragma omp parallel for schedule(dynamic,256) reduction(+:nn)
  for (int64_t k = 0; k < nel; k++) {
    int64_t e = els ? els[k] : k;
    double gcoef[ORC_MAXND*3], ls[4];
    gather_iso(dim, el2v, vert, lset_v, ndg, el2dof_def, defo, e, gcoef, ls);
    int r = orc_elmat_h1_iso(dim, fe_order, dord, gcoef, ls, dt, intorder, form, elmats + (size_t)k*nd*nd);
    if (npts) npts[k] = r;
    if (r < 0) memset(elmats + (size_t)k*nd*nd, 0, sizeof(double)*nd*nd); else nn++;
  }
  return nn;
}

/* ------------------------------------------------------------------ quadrilaterals (groundwork; 2D) */
/* LevelsetCutQuadrilateral (cutint/straightcutrule.cpp:206-501), policy FIND_OPTIMAL, on the reference square for a bilinear level
 * set given by its vertex values at (0,0), (1,0), (1,1), (0,1).  C twin of oracle/quad_ref.py (which also covers hexahedra); both
 * are pinned against tests/golden/quad_hex_rules.npz (rules of the compiled reference). */
typedef struct { double c00, c10, c01, c11; } qls_t;          /* LevelsetWrapper::GetCoeffsFromVals ET_QUAD (:512-514) */
static double qls(const qls_t *c, double x, double y) { return c->c00 + c->c10*x + c->c01*y + c->c11*x*y; }
static int qcheck(int n, const double *v, double eps)         /* CheckIfStraightCut (:11-39) */
{
  int hp = 0, hn = 0;
  for (int i = 0; i < n; i++) { if (v[i] > eps) hp = 1; if (v[i] < -eps) hn = 1; }
  if (hp == hn) return IF;
  return hn ? NEG : POS;
}
typedef struct { double *p, *w; int n, cap, err; } qout_t;
static void qpush(qout_t *o, double x, double y, double w)
{
  if (o->n >= o->cap) { o->err = ORC_CAP; return; }
  o->p[3*o->n] = x; o->p[3*o->n + 1] = y; o->p[3*o->n + 2] = 0.0; o->w[o->n] = w; o->n++;
}
/* Quadrilateral::GetPlainIntegrationRule (:102-128): tensor Gauss rule, x outer, y inner */
static void qplain(double x0, double x1, double y0, double y1, int order, qout_t *o)
{
  const double (*r)[4];
  int n = select_rule(1, order, &r);
  double vol = fabs((x1 - x0) * (y1 - y0));
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) qpush(o, x0 + (x1 - x0)*r[i][0], y0 + (y1 - y0)*r[j][0], r[i][3]*r[j][3]*vol);
}
/* GetTensorProductAlongXiIntegrationRule (:255-303), D = 2 */
static void qtensor(const qls_t *c, int dt, double x0, double x1, double y0, double y1, int order, qout_t *o)
{
  const double (*r)[4];
  int n = select_rule(1, order, &r);
  for (int i = 0; i < n; i++) {
    double ya = y0 + r[i][0]*(y1 - y0);
    double proj[2] = { qls(c, x1, ya), qls(c, x0, ya) };          /* ET_SEGM vertex 0 sits at xhat = 1 */
    if (qcheck(2, proj, 0.0) != IF || proj[0] == 0.0 || proj[1] == 0.0) { o->err = ORC_THROW; return; }
    double p2[3*64], w2[64];
    int k = orc_cut_rule_untrafo(1, proj, dt, order, p2, w2, 64);
    for (int q = 0; q < k; q++) {
      double x = x0 + p2[3*q]*(x1 - x0), sc = 1.0;
      if (dt == IF) {
        double gx = c->c10 + c->c11*ya, gy = c->c01 + c->c11*x;
        sc = sqrt(gx*gx + gy*gy) / fabs(gx);
      }
      qpush(o, x, ya, w2[q]*r[i][3]*(y1 - y0)*sc);
    }
  }
}
/* GetIntegrationRuleAlongXi (:453-468) with HasTopologyChangeAlongXi (:206-218) and Decompose (:220-251) on the unit square */
static void qalong(const qls_t *c, int dt, int order, qout_t *o)
{
  double v[4] = { qls(c, 0, 0), qls(c, 1, 0), qls(c, 1, 1), qls(c, 0, 1) };
  int d = qcheck(4, v, 0.0);
  if (d != IF) { if (d == dt) qplain(0, 1, 0, 1, order, o); return; }
  double xs[4] = { 0.0, 1.0, 0, 0 };
  int nx = 2;
  const int ea[2] = { 1, 0 }, eb[2] = { 2, 3 };
  for (int e = 0; e < 2; e++) {
    double ev[2] = { v[eb[e]], v[ea[e]] };
    if (qcheck(2, ev, 0.0) == IF) {
      double xi = 1.0 + ev[0] / (ev[0] - ev[1]) * (0.0 - 1.0);
      int dup = 0;
      for (int k = 0; k < nx; k++) if (xs[k] == xi) dup = 1;
      if (!dup) xs[nx++] = xi;
    }
  }
  if (nx == 2) { qtensor(c, dt, 0, 1, 0, 1, order, o); return; }
  for (int a = 0; a < nx; a++) for (int b = a + 1; b < nx; b++) if (xs[b] < xs[a]) { double t = xs[a]; xs[a] = xs[b]; xs[b] = t; }
  for (int k = 0; k + 1 < nx; k++) {
    double sv[4] = { qls(c, 0, xs[k]), qls(c, 1, xs[k]), qls(c, 1, xs[k+1]), qls(c, 0, xs[k+1]) };
    int d2 = qcheck(4, sv, 1e-15);
    if (d2 == IF) qtensor(c, dt, 0, 1, xs[k], xs[k+1], order, o);
    else if (d2 == dt) qplain(0, 1, xs[k], xs[k+1], order, o);
  }
}
/* GetFallbackIntegrationRule (:470-489): triangles {0,1,3}, {2,1,3}, level set re-interpolated at their vertices (clamped) */
static void qfallback(const qls_t *c, int dt, int order, qout_t *o)
{
  static const double CR[4][2] = { {0,0}, {1,0}, {1,1}, {0,1} };
  static const int TRI[2][3] = { {0,1,3}, {2,1,3} };
  for (int t = 0; t < 2; t++) {
    double P[3][2], v[3];
    for (int k = 0; k < 3; k++) {
      P[k][0] = CR[TRI[t][k]][0]; P[k][1] = CR[TRI[t][k]][1];
      double d = qls(c, P[k][0], P[k][1]);
      v[k] = fabs(d) > 1e-14 ? d : 1e-14;                           /* LevelsetWrapper::update_initial_coefs (:547-556) */
    }
    int d = qcheck(3, v, 0.0);
    if (d != IF && d == dt) {                                       /* SimpleX::GetPlainIntegrationRule (:80-100), base point p0 */
      const double (*r)[4];
      int n = select_rule(2, order, &r);
      double det = fabs((P[1][0]-P[0][0])*(P[2][1]-P[0][1]) - (P[2][0]-P[0][0])*(P[1][1]-P[0][1]));
      for (int q = 0; q < n; q++) {
        double l0 = 1.0 - r[q][0] - r[q][1];
        qpush(o, l0*P[0][0] + r[q][0]*P[1][0] + r[q][1]*P[2][0], l0*P[0][1] + r[q][0]*P[1][1] + r[q][1]*P[2][1], r[q][3]*det);
      }
    } else if (d == IF && dt != IF) {                               /* volumes: affine image of the reference-triangle cut rule */
      double p2[3*256], w2[256];
      int n = orc_cut_rule_untrafo(2, v, dt, order, p2, w2, 256);
      if (n == ORC_THROW) { o->err = ORC_THROW; return; }
      double det = fabs((P[0][0]-P[2][0])*(P[1][1]-P[2][1]) - (P[1][0]-P[2][0])*(P[0][1]-P[2][1]));
      for (int q = 0; q < n; q++) {
        double a = p2[3*q], b = p2[3*q+1], l2 = 1.0 - a - b;
        qpush(o, a*P[0][0] + b*P[1][0] + l2*P[2][0], a*P[0][1] + b*P[1][1] + l2*P[2][1], w2[q]*det);
      }
    } else if (d == IF) {                                           /* interface: one segment between the two edge zeros */
      double cut[3][2];
      int nc = 0;
      for (int i = 0; i < 3; i++)
        for (int j = i + 1; j < 3; j++)
          if ((v[i] >= 0) != (v[j] >= 0)) {
            double f = v[i] / (v[i] - v[j]);
            cut[nc][0] = P[i][0] + f*(P[j][0] - P[i][0]); cut[nc][1] = P[i][1] + f*(P[j][1] - P[i][1]); nc++;
          }
      if (nc != 2) { o->err = ORC_THROW; return; }
      const double (*r)[4];
      int n = select_rule(1, order, &r);
      double len = sqrt((cut[1][0]-cut[0][0])*(cut[1][0]-cut[0][0]) + (cut[1][1]-cut[0][1])*(cut[1][1]-cut[0][1]));
      for (int q = 0; q < n; q++)
        qpush(o, (1.0 - r[q][0])*cut[0][0] + r[q][0]*cut[1][0], (1.0 - r[q][0])*cut[0][1] + r[q][0]*cut[1][1], r[q][3]*len);
    }
  }
}
/* StraightCutIntegrationRuleUntransformed for ET_QUAD (:663-726) + GetIntegrationRule / GetDimensionSwap (:415-434, :491-500) */
int orc_quad_rule_untrafo(const double *vals, int dt, int order, double *pts3, double *wts, int cap)
{
  qout_t o = { pts3, wts, 0, cap, 0 };
  int d = qcheck(4, vals, 0.0);
  if (d != IF) {
    if (d != dt) return ORC_NULL;
    qplain(0, 1, 0, 1, order, &o);
    return o.err ? o.err : o.n;
  }
  qls_t c;
  c.c00 = vals[0]; c.c10 = vals[1] - c.c00; c.c01 = vals[3] - c.c00; c.c11 = vals[2] - c.c10 - c.c01 - c.c00;
  /* GetExactCritsQBound2D (:392-413) */
  int allow[2] = { 1, 1 };
  double h = -c.c10 / c.c11;
  if (h > 0 && h < 1) allow[1] = 0;
  h = -c.c01 / c.c11;
  if (h > 0 && h < 1) allow[0] = 0;
  double qm[2] = { 0, 0 };
  static const double CR[4][2] = { {0,0}, {1,0}, {1,1}, {0,1} };
  for (int k = 0; k < 4; k++) {
    double gx = c.c10 + c.c11*CR[k][1], gy = c.c01 + c.c11*CR[k][0], nr = sqrt(gx*gx + gy*gy);
    double qy = fabs(gy)/nr, qx = fabs(gx)/nr;
    if (qy > qm[1]) qm[1] = qy;
    if (qx > qm[0]) qm[0] = qx;
  }
  for (int k = 0; k < 2; k++) if (!allow[k]) qm[k] = 2.0;
  const double cb = 0.999;
  int sw;                                                           /* 0 ID, 1 X_Y, 2 NONE */
  if (qm[0] < cb && qm[1] < cb) sw = qm[1] <= qm[0] ? 0 : 1;
  else if (qm[1] < cb) sw = 0;
  else if (qm[0] < cb) sw = 1;
  else sw = 2;
  if (sw == 0) qalong(&c, dt, order, &o);
  else if (sw == 1) {                                               /* GetIntegrationRuleOnXYPermutatedQuad (:305-320) */
    qls_t cr = { c.c00, c.c01, c.c10, c.c11 };
    qalong(&cr, dt, order, &o);
    for (int k = 0; k < o.n; k++) { double t = pts3[3*k]; pts3[3*k] = pts3[3*k+1]; pts3[3*k+1] = t; }
  } else qfallback(&c, dt, order, &o);
  return o.err ? o.err : o.n;
}

/* ------------------------------------------------------------------ mesh-level loops */
static void gather_el(int dim, const int32_t *el2v, const double *vert, const double *lset_v, int64_t e, double *pmat, double *ls, int *vn)
{
  for (int i = 0; i < dim + 1; i++) {
    int v = el2v[e*(dim+1) + i];
    for (int d = 0; d < dim; d++) pmat[i*dim + d] = vert[(int64_t)v*dim + d];
    ls[i] = lset_v[v];
    if (vn) vn[i] = v;
  }
}

/* utils/p1interpol.cpp:86-90 clamp */
void orc_interpolate_p1_clamp(int64_t nv, const double *in, double eps, double *out)
{
  for (int64_t i = 0; i < nv; i++) { double v = in[i]; if (fabs(v) < eps) v = eps; out[i] = v; }
}

/* xfem/cutinfo.cpp:54-120 CutInformation::Update, VOL part. dom[e] in {NEG,POS,IF} or -2 where the reference throws. returns #throws */
int64_t orc_cutinfo_mesh(int dim, int64_t ne, const int32_t *el2v, const double *vert, const double *lset_v,
                         int8_t *dom, double *ratio)
{
  int64_t nerr = 0;
#pragma omp parallel for schedule(static) reduction(+:nerr)
  for (int64_t e = 0; e < ne; e++) {
    double pmat[12], ls[4];
    gather_el(dim, el2v, vert, lset_v, e, pmat, ls, 0);
    int r = orc_cutinfo_element(dim, pmat, ls, &ratio[e]);
    dom[e] = (int8_t)r;
    if (r < 0) nerr++;
  }
  return nerr;
}

/* per element rule sizes for (dt, order): -1 = no rule. */
void orc_rule_count_mesh(int dim, int64_t ne, const int32_t *el2v, const double *lset_v, int dt, int order, int32_t *npts)
{
#pragma omp parallel for schedule(static)
  for (int64_t e = 0; e < ne; e++) {
    static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS];
    double ls[4];
    for (int i = 0; i < dim + 1; i++) ls[i] = lset_v[el2v[e*(dim+1) + i]];
    npts[e] = orc_cut_rule_untrafo(dim, ls, dt, order, pts, w, ORC_MAXPTS);
  }
}

/* cutint/python_cutint.cpp:41-68 IntegrateX, cf = polynomial; per-element partials summed in element order (the reference's
 * AtomicAdd order is thread-dependent; any order is "the reference" up to re-association). */
double orc_integrate_mesh(int dim, int64_t ne, const int32_t *el2v, const double *vert, const double *lset_v, int dt, int order,
                          int nmono, const double *mono, int64_t *npts_total)
{
  double sum = 0;
  int64_t nptot = 0;
#pragma omp parallel for schedule(static) reduction(+:sum,nptot)
  for (int64_t e = 0; e < ne; e++) {
    static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS];
    double pmat[12], ls[4], x[3];
    gather_el(dim, el2v, vert, lset_v, e, pmat, ls, 0);
    int np = orc_cut_rule(dim, pmat, ls, dt, order, pts, w, ORC_MAXPTS);
    if (np < 0) continue;
    affine_t A;
    affine_init(dim, pmat, &A);
    double meas = fabs(A.det), lsum = 0;
    for (int q = 0; q < np; q++) {
      affine_map(&A, pts + 3*q, x);
      lsum += meas * w[q] * eval_mono(nmono, mono, x);
    }
    sum += lsum;
    nptot += np;
  }
  if (npts_total) *npts_total = nptot;
  return sum;
}

/* Space-time element matrix: T_CalcElementMatrixAdd (xfem/symboliccutbfi.cpp:73-315) with the space-time rule and the tensor-product
 * element SpaceTimeFE (spacetime/SpaceTimeFE.cpp:29-113): shape[j*nd_s + i] = phi_i(x) tau_j(tref) (CalcShape :29-55), spatial
 * gradient grad phi_i tau_j (CalcDShape :57-84), reference-time derivative phi_i tau_j' (CalcDtShape :88-113); tau = NodalTimeFE of
 * order ntd_fe - 1.  elmat[(test)][(trial)] = sum_q |det J| w_q (c_m u v + c_s grad u . grad v + c_t dt(u) v).  lset_st [ntd][dim+1]. */
int orc_st_elmat_h1(int dim, int fe_order, int ntd_fe, const double *pmat, const double *lset_st, int ntd, int dt, int order_time,
                    int order_space, double c_mass, double c_stiff, double c_dt, double *elmat)
{
  static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS], tq[ORC_MAXPTS];
  if (ntd_fe < 1 || ntd_fe > ST_MAXTD) return ORC_UNSUP;
  int np = orc_st_cut_rule(dim, pmat, lset_st, ntd, dt, order_time, order_space, pts, tq, w, ORC_MAXPTS);
  int nds = orc_h1_ndof(dim, fe_order), N = nds * ntd_fe;
  for (int i = 0; i < N*N; i++) elmat[i] = 0;
  if (np < 0) return np;
  affine_t A;
  affine_init(dim, pmat, &A);
  double meas = fabs(A.det);
  double shape[ORC_MAXND], dref[ORC_MAXND*3], dphys[ORC_MAXND*3], tau[ST_MAXTD], dtau[ST_MAXTD];
  for (int q = 0; q < np; q++) {
    if (orc_h1_shape(dim, fe_order, NULL, pts + 3*q, shape, dref) < 0) return ORC_UNSUP;
    for (int i = 0; i < nds; i++)
      for (int d = 0; d < dim; d++) {
        double v = 0;
        for (int k = 0; k < dim; k++) v += dref[i*dim + k] * A.Jinv[k][d];
        dphys[i*dim + d] = v;
      }
    st_lagrange(ntd_fe, tq[q], tau, dtau);
    double fac = meas * w[q];
    for (int j = 0; j < ntd_fe; j++) for (int i = 0; i < nds; i++)
      for (int jp = 0; jp < ntd_fe; jp++) for (int ip = 0; ip < nds; ip++) {
        double g = 0;
        for (int d = 0; d < dim; d++) g += dphys[i*dim + d] * dphys[ip*dim + d];
        double v = (c_mass * shape[i] * shape[ip] + c_stiff * g) * tau[j] * tau[jp] + c_dt * shape[i] * tau[j] * shape[ip] * dtau[jp];
        elmat[(j*nds + i)*N + jp*nds + ip] += fac * v;
      }
  }
  return np;
}

/* xfem/cutinfo.cpp:79-111 with time_order >= 0: part volumes of the space-time prism from the POS and NEG rules of order 0 in space */
int orc_st_cutinfo_element(int dim, const double *pmat, const double *lset_st, int ntd, int time_order, double *ratio)
{
  static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS], tq[ORC_MAXPTS];
  double pv[2] = {0, 0};
  const int order_np[2] = {POS, NEG};
  for (int k = 0; k < 2; k++) {
    int np = order_np[k];
    int n = orc_st_cut_rule(dim, pmat, lset_st, ntd, np, time_order, 0, pts, tq, w, ORC_MAXPTS);
    if (n == ORC_THROW || n == ORC_UNSUP || n == ORC_CAP) return n;
    for (int i = 0; i < n; i++) pv[np] += w[i];
  }
  *ratio = pv[NEG] / (pv[NEG] + pv[POS]);
  if (pv[NEG] > 0.0) return pv[POS] > 0.0 ? IF : NEG;
  return POS;
}

/* IntegrateX with time_order >= 0 (cutint/python_cutint.cpp:20-72 -> CreateCutIntegrationRule, cutint/xintegration.cpp:38-58):
 * sum_e sum_q |det J_e| w_q f(x_q, t_q); mono rows {coef, ex, ey, ez, et} (et = power of the reference time tref).
 * lset_st [ntd][nvtot] global space-time vector (dof = t * ndof_space + vertex, spacetime/SpaceTimeFESpace.cpp:109-122). */
double orc_st_integrate_mesh(int dim, int64_t ne, const int32_t *el2v, const double *vert, int64_t nvtot, const double *lset_st, int ntd,
                             int dt, int order_space, int order_time, int nmono, const double *mono5, int64_t *npts_total)
{
  double sum = 0;
  int64_t nptot = 0;
  const int nv = dim + 1;
#pragma omp parallel for schedule(static) reduction(+:sum,nptot)
  for (int64_t e = 0; e < ne; e++) {
    static _Thread_local double pts[3*ORC_MAXPTS], w[ORC_MAXPTS], tq[ORC_MAXPTS];
    double pmat[12], ls0[4], x[3], lst[4*ST_MAXTD];
    gather_el(dim, el2v, vert, lset_st, e, pmat, ls0, 0);
    for (int k = 0; k < ntd; k++)
      for (int v = 0; v < nv; v++) lst[k*nv + v] = lset_st[k*nvtot + el2v[e*nv + v]];
    int np = orc_st_cut_rule(dim, pmat, lst, ntd, dt, order_time, order_space, pts, tq, w, ORC_MAXPTS);
    if (np < 0) continue;
    affine_t A;
    affine_init(dim, pmat, &A);
    double meas = fabs(A.det), lsum = 0;
    for (int q = 0; q < np; q++) {
      affine_map(&A, pts + 3*q, x);
      double f = 0;
      for (int m = 0; m < nmono; m++)
        f += mono5[5*m] * pow(x[0], mono5[5*m+1]) * pow(x[1], mono5[5*m+2]) * pow(x[2], mono5[5*m+3]) * pow(tq[q], mono5[5*m+4]);
      lsum += meas * w[q] * f;
    }
    sum += lsum;
    nptot += np;
  }
  if (npts_total) *npts_total = nptot;
  return sum;
}

/* batched element matrices (els == NULL: all elements) */
int64_t orc_elmat_mesh(int dim, int64_t nel, const int64_t *els, const int32_t *el2v, const double *vert, const double *lset_v,
                       int fe_order, int dt, int intorder, double c_mass, double c_stiff, double *elmats, int32_t *npts)
{
  int nd = orc_h1_ndof(dim, fe_order);
  int64_t nn = 0;
#pragma omp parallel for schedule(dynamic,256) reduction(+:nn)
  for (int64_t k = 0; k < nel; k++) {
    int64_t e = els ? els[k] : k;
    double pmat[12], ls[4];
    int vn[4];
    gather_el(dim, el2v, vert, lset_v, e, pmat, ls, vn);
    int r = orc_elmat_h1(dim, fe_order, vn, pmat, ls, dt, intorder, c_mass, c_stiff, elmats + (size_t)k*nd*nd);
    if (npts) npts[k] = r;
    if (r < 0) memset(elmats + (size_t)k*nd*nd, 0, sizeof(double)*nd*nd); else nn++;
  }
  return nn;
}

/* NGS/comp/bilinearform.cpp:1117-1563 element loop (colour by colour, NGS/comp/fespace.cpp:1540-1566) +
 * NGS/linalg/sparsematrix_impl.hpp:63-103 SparseMatrixTM::AddElementMatrix (sorted columns, linear scan of the row). */
int64_t orc_assemble_csr(int dim, const int32_t *el2v, const double *vert, const double *lset_v,
                         int fe_order, const int32_t *el2dof, int dt, int intorder, double c_mass, double c_stiff,
                         int ncolors, const int64_t *color_ptr, const int32_t *color_els,
                         const int64_t *rowptr, const int32_t *colidx, double *vals)
{
  int nd = orc_h1_ndof(dim, fe_order);
  int64_t nass = 0;
  for (int c = 0; c < ncolors; c++) {
#pragma omp parallel for schedule(dynamic,64) reduction(+:nass)
    for (int64_t k = color_ptr[c]; k < color_ptr[c+1]; k++) {
      int64_t e = color_els[k];
      double pmat[12], ls[4], elmat[ORC_MAXND*ORC_MAXND];
      int vn[4];
      gather_el(dim, el2v, vert, lset_v, e, pmat, ls, vn);
      int r = orc_elmat_h1(dim, fe_order, vn, pmat, ls, dt, intorder, c_mass, c_stiff, elmat);
      if (r < 0) continue;
      nass++;
      const int32_t *dn = el2dof + e*nd;
      int map[ORC_MAXND];
      for (int i = 0; i < nd; i++) map[i] = i;
      for (int i = 1; i < nd; i++) {                  /* QuickSortI(dnums2, map) */
        int m = map[i], j = i;
        while (j > 0 && dn[map[j-1]] > dn[m]) { map[j] = map[j-1]; j--; }
        map[j] = m;
      }
      for (int i = 0; i < nd; i++) {
        int64_t k0 = rowptr[dn[i]];
        for (int j1 = 0; j1 < nd; j1++) {
          int j = map[j1];
          while (colidx[k0] != dn[j]) k0++;
          vals[k0] += elmat[i*nd + j];
        }
      }
    }
  }
  return nass;
}

/* NGS/comp/linearform.cpp:209-243 + symboliccutlfi: vec[dof] += elvec (serial, element order) */
int64_t orc_assemble_vec(int dim, int64_t ne, const int32_t *el2v, const double *vert, const double *lset_v,
                         int fe_order, const int32_t *el2dof, int dt, int intorder, int nmono, const double *mono, double *vec)
{
  int nd = orc_h1_ndof(dim, fe_order);
  int64_t nass = 0;
  for (int64_t e = 0; e < ne; e++) {
    double pmat[12], ls[4], elvec[ORC_MAXND];
    int vn[4];
    gather_el(dim, el2v, vert, lset_v, e, pmat, ls, vn);
    int r = orc_elvec_h1(dim, fe_order, vn, pmat, ls, dt, intorder, nmono, mono, elvec);
    if (r < 0) continue;
    nass++;
    for (int i = 0; i < nd; i++) vec[el2dof[e*nd + i]] += elvec[i];
  }
  return nass;
}

/* xfem/cutinfo.cpp:191-255 GetFacetsWithNeighborTypes (no periodic identification).  facet2el[f][2], -1 = no neighbour.
 * a,b,out are byte masks (0/1) here; the product packs them as NGSolve BitArray words. */
void orc_facets_neighbor_types(int64_t nf, const int32_t *facet2el, const uint8_t *a, const uint8_t *b,
                               int bnd_a, int bnd_b, int use_and, uint8_t *out)
{
  for (int64_t f = 0; f < nf; f++) {
    int e0 = facet2el[2*f], e1 = facet2el[2*f + 1];
    out[f] = 0;
    if (e0 < 0) continue;
    int al = a[e0], ar = e1 >= 0 ? a[e1] : bnd_a;
    int bl = b[e0], br = e1 >= 0 ? b[e1] : bnd_b;
    if (use_and) out[f] = (al && br) || (ar && bl);
    else out[f] = (al || br) || (ar || bl);
  }
}

/* xfem/cutinfo.cpp:257-279 */
void orc_elements_with_neighbor_facets(int64_t nf, int64_t ne, const int32_t *facet2el, const uint8_t *a, uint8_t *out)
{
  memset(out, 0, ne);
  for (int64_t f = 0; f < nf; f++)
    if (a[f]) for (int k = 0; k < 2; k++) if (facet2el[2*f + k] >= 0) out[facet2el[2*f + k]] = 1;
}

/* xfem/cutinfo.cpp:281-304 */
void orc_dofs_of_elements(int64_t ne, int nd, int64_t ndof, const int32_t *el2dof, const uint8_t *a, uint8_t *out)
{
  memset(out, 0, ndof);
  for (int64_t e = 0; e < ne; e++)
    if (a[e]) for (int i = 0; i < nd; i++) out[el2dof[e*nd + i]] = 1;
}

/* xfem/symboliccutbfi.cpp:1202-1411 SymbolicFacetPatchBilinearFormIntegrator::CalcFacetMatrix for the ghost-penalty form
 *   c * h^hpow * (u - u.Other()) * (v - v.Other())       (py_tutorials/fictdom_ghostpen.py:105-107)
 * on affine simplices: volume rules of order 2*fe_order on BOTH elements (:1219-1220); each point is pulled back into
 * the other element (MapPatchIntegrationPoint :1115-1199 -- exact in one step for affine maps); weight mir.GetWeight().
 * elmat is (2 nd)x(2 nd) row-major over [dofs of el1 | dofs of el2]. */
int orc_facetpatch_elmat_h1(int dim, int fe_order, const int *vn1, const double *pmat1, const int *vn2, const double *pmat2,
                            double coef, int hpow, double *elmat)
{
  int nd = orc_h1_ndof(dim, fe_order), n2 = 2*nd;
  const double (*r)[4];
  int nr = select_rule(dim, 2*fe_order, &r);
  if (nr < 0) return nr;
  affine_t A[2];
  affine_init(dim, pmat1, &A[0]);
  affine_init(dim, pmat2, &A[1]);
  for (int i = 0; i < n2*n2; i++) elmat[i] = 0;
  /* the coefficient (h = specialcf.mesh_size, NGS/fem/python_fem.cpp:564-593) is evaluated on mir1, i.e. on element 1,
   * for all patch points (:1359) */
  coef *= pow(pow(fabs(A[0].det), 1.0 / dim), hpow);
  double sh[2][ORC_MAXND], dref[ORC_MAXND*3], jump[2*ORC_MAXND];
  for (int side = 0; side < 2; side++) {
    const affine_t *Ah = &A[side], *Ao = &A[1-side];
    const int *vh = side == 0 ? vn1 : vn2, *vo = side == 0 ? vn2 : vn1;
    double meas = fabs(Ah->det);
    for (int q = 0; q < nr; q++) {
      double xi[3] = {r[q][0], r[q][1], r[q][2]}, x[3], xo[3] = {0, 0, 0};
      affine_map(Ah, xi, x);
      for (int k = 0; k < dim; k++) {                  /* xo = Finv_other (x - p0_other) */
        double v = 0;
        for (int rr = 0; rr < dim; rr++) v += Ao->Jinv[k][rr] * (x[rr] - Ao->p0[rr]);
        xo[k] = v;
      }
      orc_h1_shape(dim, fe_order, vh, xi, sh[0], dref);
      orc_h1_shape(dim, fe_order, vo, xo, sh[1], dref);
      /* u - u.Other(): +shape on the element owning the point, -shape (extended polynomial) on the other */
      for (int i = 0; i < nd; i++) {
        jump[(side == 0 ? 0 : nd) + i] = sh[0][i];
        jump[(side == 0 ? nd : 0) + i] = -sh[1][i];
      }
      double fac = coef * meas * r[q][3];
      for (int i = 0; i < n2; i++)
        for (int j = 0; j < n2; j++) elmat[i*n2 + j] += fac * jump[i] * jump[j];
    }
  }
  return 2*nr;
}

/* MapPatchIntegrationPoint (xfem/symboliccutbfi.cpp:1113-1199) through isoparametric maps: the point xi_from of the `from`
 * element is pulled into the `to` element by Newton's method (start: the affine map through the reference vertices, :1141-1166;
 * at most 200 steps; the loop stops when the residual BEFORE the last update was <= 1e-12 sqrt(det F_from), :1168-1180 -- with a
 * negative Jacobian determinant the bound is NaN and exactly one step is taken).  Returns the weight the reference stores:
 * from_weight / measure_to, measure_to = |det F_to| at the point of the last evaluation (:1179, 1193). */
static double map_patch_point(int dim, int dord, const double *g_from, const double *g_to, const double *xi_from, double w_ref,
                              double *xi_to)
{
  double x[3], J[3][3], Ji[3][3], det;
  iso_eval(dim, dord, g_from, xi_from, x, J, Ji, &det);
  const double h = sqrt(det);
  const double wfrom = fabs(det) * w_ref;                       /* mip.GetWeight() */
  double z[3] = {0, 0, 0}, x0[3], A[3][3], Ai[3][3], f[3], pm[12], t[3][3], d0;
  iso_eval(dim, dord, g_to, z, x0, t, Ai, &d0);
  for (int d = 0; d < dim; d++) {
    double e[3] = {0, 0, 0}, xe[3];
    e[d] = 1.0;
    iso_eval(dim, dord, g_to, e, xe, t, Ai, &d0);
    for (int r = 0; r < dim; r++) A[r][d] = xe[r] - x0[r];
  }
  for (int k = 0; k < dim; k++) for (int r = 0; r < dim; r++) pm[k*dim + r] = A[r][k];
  for (int r = 0; r < dim; r++) pm[dim*dim + r] = 0.0;
  affine_t AA;
  affine_init(dim, pm, &AA);
  const double w00 = fabs(AA.det);
  double xi00[3] = {0, 0, 0}, xi[3] = {0, 0, 0};
  for (int r = 0; r < dim; r++) f[r] = x[r] - x0[r];
  for (int k = 0; k < dim; k++) { double v = 0; for (int r = 0; r < dim; r++) v += AA.Jinv[k][r] * f[r]; xi00[k] = v; xi[k] = v; }
  int its = 0;
  double w = 0, nrm = 0;
  while (its == 0 || (nrm > 1e-12 * h && its < 200)) {
    double xq[3], Jq[3][3], Jqi[3][3], dq, diff[3];
    iso_eval(dim, dord, g_to, xi, xq, Jq, Jqi, &dq);
    nrm = 0;
    for (int r = 0; r < dim; r++) { diff[r] = x[r] - xq[r]; nrm += diff[r]*diff[r]; }
    nrm = sqrt(nrm);
    for (int k = 0; k < dim; k++) { double v = 0; for (int r = 0; r < dim; r++) v += Jqi[k][r] * diff[r]; xi[k] += v; }
    its++;
    w = fabs(dq);
  }
  if (its >= 200) { for (int k = 0; k < 3; k++) xi_to[k] = xi00[k]; return wfrom / w00; }
  for (int k = 0; k < 3; k++) xi_to[k] = xi[k];
  return wfrom / w;
}

/* SymbolicFacetPatchBilinearFormIntegrator::CalcFacetMatrix (:1202-1411) on a DEFORMED mesh (isoparametric elements of order
 * dord, geometry coefficients g1 / g2), form coef h^hpow (u - u.Other())(v - v.Other()); h = specialcf.mesh_size on mir1 per point.
 * Pinned against ref_facetpatch_elmat_h1 (the reference's own MapPatchIntegrationPoint). */
int orc_facetpatch_elmat_h1_iso(int dim, int fe_order, int dord, const int *vn1, const double *g1, const int *vn2, const double *g2,
                                double coef, int hpow, double *elmat)
{
  int nd = orc_h1_ndof(dim, fe_order), n2 = 2*nd;
  const double (*r)[4];
  int nr = select_rule(dim, 2*fe_order, &r);
  if (nr < 0) return nr;
  for (int i = 0; i < n2*n2; i++) elmat[i] = 0;
  double sh1[ORC_MAXND], sh2[ORC_MAXND], dref[ORC_MAXND*3], jump[2*ORC_MAXND];
  for (int l = 0; l < 2*nr; l++) {
    double xi1[3] = {0, 0, 0}, xi2[3] = {0, 0, 0}, w1;
    if (l < nr) {
      for (int k = 0; k < 3; k++) xi1[k] = r[l][k];
      map_patch_point(dim, dord, g1, g2, xi1, r[l][3], xi2);
      w1 = r[l][3];
    } else {
      for (int k = 0; k < 3; k++) xi2[k] = r[l - nr][k];
      w1 = map_patch_point(dim, dord, g2, g1, xi2, r[l - nr][3], xi1);
    }
    double x[3], J[3][3], Ji[3][3], det;
    iso_eval(dim, dord, g1, xi1, x, J, Ji, &det);                 /* mir1[l] */
    double hh = pow(fabs(det), 1.0 / dim);
    double val = coef * pow(hh, hpow) * (fabs(det) * w1);
    orc_h1_shape(dim, fe_order, vn1, xi1, sh1, dref);
    orc_h1_shape(dim, fe_order, vn2, xi2, sh2, dref);
    for (int i = 0; i < nd; i++) { jump[i] = sh1[i]; jump[nd + i] = -sh2[i]; }
    for (int i = 0; i < n2; i++)
      for (int j = 0; j < n2; j++) elmat[i*n2 + j] += val * jump[i] * jump[j];
  }
  return 2*nr;
}

int orc_max_threads(void)
{
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void orc_set_threads(int n)
{
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* ------------------------------------------------------------------ set-up helpers of the CPU arm (bench.py --impl reference)
 * The reference arm must run on a GPU-less host without the product library: sparsity pattern and element colouring are
 * built here.  Both follow NGSolve's constructions in spirit (MatrixGraph from the element -> dof table,
 * NGS/linalg/sparsematrix.cpp; greedy element colouring, NGS/comp/fespace.cpp:598-705) -- any valid pattern / colouring
 * gives the same CSR values up to summation order. */

/* pattern of the listed elements' couplings: rowptr[ndof+1]; colidx == NULL: count only.  table [n][nd] = dofs of list entry k */
int64_t orc_csr_pattern(int64_t n, int nd, const int32_t *table, int64_t ndof, int64_t *rowptr, int32_t *colidx)
{
  /* dof -> list entries (transposed table) */
  int64_t *cnt = (int64_t *)calloc((size_t)ndof + 1, sizeof(int64_t));
  for (int64_t p = 0; p < n*nd; p++) cnt[table[p] + 1]++;
  for (int64_t d = 0; d < ndof; d++) cnt[d + 1] += cnt[d];
  int32_t *d2e = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n*nd > 0 ? n*nd : 1));
  int64_t *fill = (int64_t *)malloc(sizeof(int64_t) * (size_t)(ndof > 0 ? ndof : 1));
  memcpy(fill, cnt, sizeof(int64_t) * (size_t)ndof);
  for (int64_t k = 0; k < n; k++) for (int i = 0; i < nd; i++) d2e[fill[table[k*nd + i]]++] = (int32_t)k;
  free(fill);
  if (!colidx) {
    rowptr[0] = 0;
#pragma omp parallel
    {
      int cap = 4096;
      int32_t *buf = (int32_t *)malloc(sizeof(int32_t) * cap);
#pragma omp for schedule(dynamic,1024)
      for (int64_t d = 0; d < ndof; d++) {
        int m = (int)(cnt[d + 1] - cnt[d]) * nd;
        if (m > cap) { cap = 2*m; buf = (int32_t *)realloc(buf, sizeof(int32_t) * cap); }
        int t = 0;
        for (int64_t q = cnt[d]; q < cnt[d + 1]; q++) for (int j = 0; j < nd; j++) buf[t++] = table[(int64_t)d2e[q]*nd + j];
        for (int a = 1; a < t; a++) { int32_t v = buf[a]; int b = a; while (b > 0 && buf[b-1] > v) { buf[b] = buf[b-1]; b--; } buf[b] = v; }
        int u = 0;
        for (int a = 0; a < t; a++) if (a == 0 || buf[a] != buf[a-1]) u++;
        rowptr[d + 1] = u;
      }
      free(buf);
    }
    for (int64_t d = 0; d < ndof; d++) rowptr[d + 1] += rowptr[d];
  } else {
#pragma omp parallel
    {
      int cap = 4096;
      int32_t *buf = (int32_t *)malloc(sizeof(int32_t) * cap);
#pragma omp for schedule(dynamic,1024)
      for (int64_t d = 0; d < ndof; d++) {
        int m = (int)(cnt[d + 1] - cnt[d]) * nd;
        if (m > cap) { cap = 2*m; buf = (int32_t *)realloc(buf, sizeof(int32_t) * cap); }
        int t = 0;
        for (int64_t q = cnt[d]; q < cnt[d + 1]; q++) for (int j = 0; j < nd; j++) buf[t++] = table[(int64_t)d2e[q]*nd + j];
        for (int a = 1; a < t; a++) { int32_t v = buf[a]; int b = a; while (b > 0 && buf[b-1] > v) { buf[b] = buf[b-1]; b--; } buf[b] = v; }
        int64_t o = rowptr[d];
        for (int a = 0; a < t; a++) if (a == 0 || buf[a] != buf[a-1]) colidx[o++] = buf[a];
      }
      free(buf);
    }
  }
  int64_t nnz = rowptr[ndof];
  free(cnt); free(d2e);
  return nnz;
}

/* greedy colouring of list entries: two entries of one colour share no dof.  perm[n] = colour-sorted positions,
 * color_ptr[ncolors+1]; returns ncolors (<= 128) or -1 */
int orc_color_elements(int64_t n, int nd, const int32_t *table, int64_t ndof, int64_t *perm, int64_t *color_ptr)
{
  uint64_t *used = (uint64_t *)calloc((size_t)ndof * 2 + 2, sizeof(uint64_t));
  uint8_t *col = (uint8_t *)malloc((size_t)(n > 0 ? n : 1));
  int ncol = 0;
  for (int64_t k = 0; k < n; k++) {
    uint64_t m0 = 0, m1 = 0;
    for (int i = 0; i < nd; i++) { m0 |= used[2*(int64_t)table[k*nd + i]]; m1 |= used[2*(int64_t)table[k*nd + i] + 1]; }
    int c;
    if (~m0) c = __builtin_ctzll(~m0);
    else if (~m1) c = 64 + __builtin_ctzll(~m1);
    else { free(used); free(col); return -1; }
    col[k] = (uint8_t)c;
    if (c + 1 > ncol) ncol = c + 1;
    for (int i = 0; i < nd; i++) {
      if (c < 64) used[2*(int64_t)table[k*nd + i]] |= 1ULL << c; else used[2*(int64_t)table[k*nd + i] + 1] |= 1ULL << (c - 64);
    }
  }
  for (int c = 0; c <= ncol; c++) color_ptr[c] = 0;
  for (int64_t k = 0; k < n; k++) color_ptr[col[k] + 1]++;
  for (int c = 0; c < ncol; c++) color_ptr[c + 1] += color_ptr[c];
  int64_t *fill = (int64_t *)malloc(sizeof(int64_t) * (size_t)(ncol + 1));
  memcpy(fill, color_ptr, sizeof(int64_t) * (size_t)(ncol + 1));
  for (int64_t k = 0; k < n; k++) perm[fill[col[k]]++] = k;
  free(fill); free(used); free(col);
  return ncol;
}
