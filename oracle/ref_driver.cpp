// TEST INFRASTRUCTURE -- not part of the product path.
//
// extern "C" driver around the reference's OWN compiled sources (ngsxfem
// cutint/straightcutrule.cpp + NGSolve fem/bla, compiled in place from
// /root/reference by oracle/build_ref.sh into oracle/_ref/libngsxref.so).
// Everything numerically relevant below is a call INTO reference code:
//   * xintegration::StraightCutIntegrationRule[Untransformed]   (cutint/straightcutrule.cpp:580-726)
//   * ngfem::SelectIntegrationRule                              (NGS/fem/intrule.cpp:2620-2659)
//   * ngfem::H1HighOrderFE<ET>::CalcShape / CalcMappedDShape    (NGS/fem/h1hofe_impl.hpp)
//   * ngfem::FE_ElementTransformation / MappedIntegrationRule   (NGS/fem/elementtransformation.cpp)
//   * xintegration::SpaceTimeCutIntegrationRule + ngfem::NodalTimeFE (cutint/spacetimecutrule.cpp:98-196, spacetime/SpaceTimeFE.cpp:116-173)
//   * ngbla::AddABt                                             (NGS/basiclinalg/ngblas.cpp)
//   * ngfem::T_DifferentialOperator<DiffOpX<D, RNEG/RPOS/RNEG_GRAD/RPOS_GRAD>>::CalcMatrix on ngfem::XFiniteElement
//                                                               (xfem/xfemdiffops.cpp:8-82, xfem/xfiniteelement.cpp:12-31)
//   * MapPatchIntegrationPoint<D>                               (xfem/symboliccutbfi.cpp:1113-1199; the template is lifted out
//     of the reference file at build time by oracle/build_ref.sh into the git-ignored oracle/_ref/mappatch_extract.hpp:
//     the rest of that file needs NGSolve's comp layer and SymbolicBilinearFormIntegrator, which do not link here)
// The only restated logic is the ~30-line B*D*B^T accumulation of
// xfem/symboliccutbfi.cpp:219-301 and :1337-1408 (the symbolic
// CoefficientFunction tree of the reference cannot be built without the full
// NGSolve comp layer; the coefficient functions of the tutorials' forms --
// constants, specialcf.mesh_size = |det F|^(1/D) per point (NGS/fem/python_fem.cpp:564-593), the level-set normal
// grad(lsetp1)/|grad(lsetp1)|, CutRatioGF -- are evaluated from the reference's own mapped points and shapes) and the
// CutInformation::Update decision of xfem/cutinfo.cpp:79-111.
#include <fem.hpp>
#include "../cutint/straightcutrule.hpp"
#include "../cutint/spacetimecutrule.hpp"
#include "../spacetime/SpaceTimeFE.hpp"
#include "../xfem/xfiniteelement.hpp"
#include "../xfem/xfemdiffops.hpp"
#include <omp.h>

using namespace ngfem;
using namespace xintegration;

namespace refpatch {
  using namespace std;
  using namespace ngfem;
#include "mappatch_extract.hpp"
}

namespace {

thread_local LocalHeap * tl_lh = nullptr;
LocalHeap & GetLH()
{
  if (!tl_lh) tl_lh = new LocalHeap(8000000, "ref_driver", true);
  return *tl_lh;
}


int NV(ELEMENT_TYPE et) { return ElementTopology::GetNVertices(et); }      // simplices: dim + 1; ET_QUAD 4, ET_HEX 8 (f-5 groundwork)

int CopyRule(const IntegrationRule * ir, double * pts3, double * wts, int cap)
{
  if (!ir) return -1;
  int n = ir->Size();
  if (n > cap) return -3;
  for (int i = 0; i < n; i++)
    {
      for (int d = 0; d < 3; d++) pts3[3*i+d] = (*ir)[i](d);
      wts[i] = (*ir)[i].Weight();
    }
  return n;
}

template <int D>
const IntegrationRule * TrafoRule(ELEMENT_TYPE et, const double * pmat, const double * lset, int dt, int order, LocalHeap & lh,
                                  FE_ElementTransformation<D,D> *& trafo_out)
{
  Matrix<> pm(D+1, D);
  for (int i = 0; i < D+1; i++) for (int d = 0; d < D; d++) pm(i,d) = pmat[i*D+d];
  auto trafo = new (lh) FE_ElementTransformation<D,D>(et, pm);
  trafo_out = trafo;
  Vector<> lv(D+1);
  for (int i = 0; i < D+1; i++) lv(i) = lset[i];
  return StraightCutIntegrationRule(lv, *trafo, DOMAIN_TYPE(dt), order, FIND_OPTIMAL, lh);
}

// isoparametric transformation: FE_ElementTransformation with the reference's H1HighOrderFE of order dord as the
// geometry element and pointmat = gcoef^T (NGS/fem/elementtransformation.hpp:199-260: SetElement + PointMatrix; CalcJacobian
// = pointmat * dshape).  This is what mesh.SetDeformation amounts to per element (base point + deformation field).
template <ELEMENT_TYPE ET>
FE_ElementTransformation<ET_trait<ET>::DIM, ET_trait<ET>::DIM> * IsoTrafo(int dord, const double * gcoef, LocalHeap & lh)
{
  constexpr int D = ET_trait<ET>::DIM;
  auto gfel = new (lh) H1HighOrderFE<ET>(dord);
  auto trafo = new (lh) FE_ElementTransformation<D,D>();
  trafo->SetElement(gfel, 0, 0);
  int ndg = gfel->GetNDof();
  auto pm = trafo->PointMatrix();
  for (int i = 0; i < ndg; i++) for (int d = 0; d < D; d++) pm(d, i) = gcoef[i*D+d];
  return trafo;
}

template <ELEMENT_TYPE ET>
int RuleIso(int dord, const double * gcoef, const double * lset, int dt, int order,
            double * pts3, double * wts, double * meas, double * xphys, int cap, LocalHeap & lh)
{
  constexpr int D = ET_trait<ET>::DIM;
  auto trafo = IsoTrafo<ET>(dord, gcoef, lh);
  Vector<> lv(D+1);
  for (int i = 0; i < D+1; i++) lv(i) = lset[i];
  const IntegrationRule * ir = StraightCutIntegrationRule(lv, *trafo, DOMAIN_TYPE(dt), order, FIND_OPTIMAL, lh);
  int n = CopyRule(ir, pts3, wts, cap);
  if (n < 0) return n;
  MappedIntegrationRule<D,D> mir(*ir, *trafo, lh);
  for (int i = 0; i < n; i++)
    {
      meas[i] = mir[i].GetMeasure();
      for (int d = 0; d < 3; d++) xphys[3*i+d] = d < D ? mir[i].GetPoint()(d) : 0.0;
    }
  return n;
}

// as ElmatWork::Elmat below, on the isoparametric transformation
template <ELEMENT_TYPE ET>
int ElmatIso(int fe_order, int dord, const double * gcoef, const double * lset, int dt, int intorder,
             double c_mass, double c_stiff, double * elmat_out, LocalHeap & lh)
{
  constexpr int D = ET_trait<ET>::DIM;
  H1HighOrderFE<ET> fel(fe_order);
  int nd = fel.GetNDof();
  auto trafo = IsoTrafo<ET>(dord, gcoef, lh);
  Vector<> lv(D+1);
  for (int i = 0; i < D+1; i++) lv(i) = lset[i];
  const IntegrationRule * ir = StraightCutIntegrationRule(lv, *trafo, DOMAIN_TYPE(dt), intorder, FIND_OPTIMAL, lh);
  if (!ir) return -1;
  int np = ir->Size();
  MappedIntegrationRule<D,D> mir(*ir, *trafo, lh);
  FlatMatrix<> elmat(nd, nd, lh); elmat = 0.0;
  constexpr int BS = 16;
  for (int pass = 0; pass < 2; pass++)
    {
      double coef = pass == 0 ? c_stiff : c_mass;
      if (coef == 0.0) continue;
      int pd = pass == 0 ? D : 1;
      for (int i0 = 0; i0 < np; i0 += BS)
        {
          HeapReset hr2(lh);
          int bs = min2(BS, np - i0);
          FlatMatrix<> bbmat(nd, bs*pd, lh), bdbmat(nd, bs*pd, lh);
          for (int j = 0; j < bs; j++)
            {
              int ii = i0 + j;
              double dval = coef * mir[ii].GetMeasure() * (*ir)[ii].Weight();
              if (pass == 0)
                {
                  FlatMatrixFixWidth<D> dshape(nd, lh);
                  fel.CalcMappedDShape(mir[ii], dshape);
                  for (int k = 0; k < nd; k++) for (int d = 0; d < D; d++)
                    { bbmat(k, j*D+d) = dshape(k,d); bdbmat(k, j*D+d) = dval * dshape(k,d); }
                }
              else
                {
                  FlatVector<> shape(nd, lh);
                  fel.CalcShape((*ir)[ii], shape);
                  for (int k = 0; k < nd; k++) { bbmat(k, j) = shape(k); bdbmat(k, j) = dval * shape(k); }
                }
            }
          AddABt(bbmat, bdbmat, elmat);
        }
    }
  for (int i = 0; i < nd; i++) for (int j = 0; j < nd; j++) elmat_out[i*nd+j] = elmat(i,j);
  return np;
}

// as ElmatWork::Elvec below, on the isoparametric transformation
template <ELEMENT_TYPE ET>
int ElvecIso(int fe_order, int dord, const double * gcoef, const double * lset, int dt, int intorder,
             int nmono, const double * mono, double * elvec_out, LocalHeap & lh)
{
  constexpr int D = ET_trait<ET>::DIM;
  H1HighOrderFE<ET> fel(fe_order);
  int nd = fel.GetNDof();
  auto trafo = IsoTrafo<ET>(dord, gcoef, lh);
  Vector<> lv(D+1);
  for (int i = 0; i < D+1; i++) lv(i) = lset[i];
  const IntegrationRule * ir = StraightCutIntegrationRule(lv, *trafo, DOMAIN_TYPE(dt), intorder, FIND_OPTIMAL, lh);
  if (!ir) return -1;
  int np = ir->Size();
  MappedIntegrationRule<D,D> mir(*ir, *trafo, lh);
  FlatVector<> elvec(nd, lh); elvec = 0.0;
  FlatVector<> shape(nd, lh);
  for (int i = 0; i < np; i++)
    {
      double x[3] = {0,0,0};
      for (int d = 0; d < D; d++) x[d] = mir[i].GetPoint()(d);
      double f = 0;
      for (int m = 0; m < nmono; m++)
        f += mono[4*m] * pow(x[0], mono[4*m+1]) * pow(x[1], mono[4*m+2]) * pow(x[2], mono[4*m+3]);
      double pv = mir[i].GetMeasure() * (*ir)[i].Weight() * f;
      fel.CalcShape((*ir)[i], shape);
      elvec += pv * shape;
    }
  for (int i = 0; i < nd; i++) elvec_out[i] = elvec(i);
  return np;
}

template <ELEMENT_TYPE ET>
struct ElmatWork
{
  enum { D = ET_trait<ET>::DIM };

  // returns npts (>=0), -1 if the reference returns no rule (nullptr), -2 on a reference exception
  static int Elmat(int fe_order, const int * vnums, const double * pmat, const double * lset, int dt, int intorder,
                   double c_mass, double c_stiff, double * elmat_out, LocalHeap & lh)
  {
    HeapReset hr(lh);
    H1HighOrderFE<ET> fel(fe_order);
    if (vnums) for (int i = 0; i < D+1; i++) fel.SetVertexNumber(i, vnums[i]);
    int nd = fel.GetNDof();
    FE_ElementTransformation<D,D> * trafo;
    const IntegrationRule * ir = TrafoRule<D>(ET, pmat, lset, dt, intorder, lh, trafo);
    if (!ir) return -1;
    int np = ir->Size();
    MappedIntegrationRule<D,D> mir(*ir, *trafo, lh);
    FlatMatrix<> elmat(nd, nd, lh); elmat = 0.0;
    // symboliccutbfi.cpp:219-227: D = cf-value * measure * weight-from-cut-rule
    // :246-301: blocks of BS=16 points, bbmat (ndof x bs*dim), bdbmat = bbmat*diag(D), elmat += bbmat*bdbmat^T (AddABt)
    constexpr int BS = 16;
    for (int pass = 0; pass < 2; pass++)   // proxy pairs: (grad u, grad v) then (u, v), as SymbolicBFI orders them for grad(u)*grad(v)+u*v
      {
        double coef = pass == 0 ? c_stiff : c_mass;
        if (coef == 0.0) continue;
        int pd = pass == 0 ? D : 1;
        for (int i0 = 0; i0 < np; i0 += BS)
          {
            HeapReset hr2(lh);
            int bs = min2(BS, np - i0);
            FlatMatrix<> bbmat(nd, bs*pd, lh), bdbmat(nd, bs*pd, lh);
            for (int j = 0; j < bs; j++)
              {
                int ii = i0 + j;
                double dval = coef * mir[ii].GetMeasure() * (*ir)[ii].Weight();
                if (pass == 0)
                  {
                    FlatMatrixFixWidth<D> dshape(nd, lh);
                    fel.CalcMappedDShape(mir[ii], dshape);
                    for (int k = 0; k < nd; k++) for (int d = 0; d < D; d++)
                      { bbmat(k, j*D+d) = dshape(k,d); bdbmat(k, j*D+d) = dval * dshape(k,d); }
                  }
                else
                  {
                    FlatVector<> shape(nd, lh);
                    fel.CalcShape((*ir)[ii], shape);
                    for (int k = 0; k < nd; k++) { bbmat(k, j) = shape(k); bdbmat(k, j) = dval * shape(k); }
                  }
              }
            AddABt(bbmat, bdbmat, elmat);
          }
      }
    for (int i = 0; i < nd; i++) for (int j = 0; j < nd; j++) elmat_out[i*nd+j] = elmat(i,j);
    return np;
  }

  // symboliccutlfi.cpp:85-153 with f = sum_m coef_m x^a y^b z^c (polynomial source), intorder passed by caller
  static int Elvec(int fe_order, const int * vnums, const double * pmat, const double * lset, int dt, int intorder,
                   int nmono, const double * mono, double * elvec_out, LocalHeap & lh)
  {
    HeapReset hr(lh);
    H1HighOrderFE<ET> fel(fe_order);
    if (vnums) for (int i = 0; i < D+1; i++) fel.SetVertexNumber(i, vnums[i]);
    int nd = fel.GetNDof();
    FE_ElementTransformation<D,D> * trafo;
    const IntegrationRule * ir = TrafoRule<D>(ET, pmat, lset, dt, intorder, lh, trafo);
    if (!ir) return -1;
    int np = ir->Size();
    MappedIntegrationRule<D,D> mir(*ir, *trafo, lh);
    FlatVector<> elvec(nd, lh); elvec = 0.0;
    FlatVector<> shape(nd, lh);
    for (int i = 0; i < np; i++)
      {
        double x[3] = {0,0,0};
        for (int d = 0; d < D; d++) x[d] = mir[i].GetPoint()(d);
        double f = 0;
        for (int m = 0; m < nmono; m++)
          f += mono[4*m] * pow(x[0], mono[4*m+1]) * pow(x[1], mono[4*m+2]) * pow(x[2], mono[4*m+3]);
        double pv = mir[i].GetMeasure() * (*ir)[i].Weight() * f;
        fel.CalcShape((*ir)[i], shape);
        elvec += pv * shape;
      }
    for (int i = 0; i < nd; i++) elvec_out[i] = elvec(i);
    return np;
  }
};

// element transformation of one element: dord == 0: affine from the vertex matrix geom [(D+1)][D]; dord >= 1: isoparametric,
// geom = coefficients of the H1HighOrderFE<ET>(dord) geometry element (IsoTrafo above)
template <ELEMENT_TYPE ET>
FE_ElementTransformation<ET_trait<ET>::DIM, ET_trait<ET>::DIM> * MakeTrafo(int dord, const double * geom, LocalHeap & lh)
{
  constexpr int D = ET_trait<ET>::DIM;
  if (dord >= 1) return IsoTrafo<ET>(dord, geom, lh);
  Matrix<> pm(D+1, D);
  for (int i = 0; i < D+1; i++) for (int d = 0; d < D; d++) pm(i,d) = geom[i*D+d];
  return new (lh) FE_ElementTransformation<D,D>(ET, pm);
}

// specialcf.mesh_size (NGS/fem/python_fem.cpp:564-593, volume point): |det F|^(1/D)
template <int D> double MeshSize(const MappedIntegrationPoint<D,D> & mip)
{
  double det = fabs(mip.GetJacobiDet());
  return D == 1 ? det : pow(det, 1.0/D);
}

// n = grad(lsetp1)/|grad(lsetp1)| at a mapped point: the gradient of the P1 GridFunction through the reference's own
// H1HighOrderFE(1)::CalcMappedDShape (what grad(gf) evaluates, NGS/comp/gridfunction.cpp via DiffOpGradient)
template <ELEMENT_TYPE ET>
Vec<ET_trait<ET>::DIM> LsetNormal(const MappedIntegrationPoint<ET_trait<ET>::DIM, ET_trait<ET>::DIM> & mip, const double * lset, LocalHeap & lh)
{
  constexpr int D = ET_trait<ET>::DIM;
  HeapReset hr(lh);
  H1HighOrderFE<ET> p1(1);
  FlatMatrixFixWidth<D> ds(D+1, lh);
  p1.CalcMappedDShape(mip, ds);
  Vec<D> g = 0.0;
  for (int i = 0; i < D+1; i++) for (int d = 0; d < D; d++) g(d) += lset[i] * ds(i,d);
  g /= L2Norm(g);
  return g;
}

// SymbolicCutBFI with the forms of py_tutorials/fictdom_ghostpen.py:86-96 / tracefem3d.py:74-88 on one element:
//   c_s h^ps grad u.grad v + c_m h^pm u v + c_n h^pn ((grad u.n) v + (grad v.n) u) + c_d h^pd (grad u.n)(grad v.n)
// form8 = {c_mass, c_stiff, c_nitsche, c_dndn, hpow_mass, hpow_stiff, hpow_nitsche, hpow_dndn}.  Rule, mapped points, measure,
// shapes and mapped gradients are the reference's; mesh_size and the normal are evaluated PER POINT (a coefficient tree that
// contains them is not elementwise constant, NGS/fem/coefficient.cpp:1349, xfem/symboliccutbfi.cpp:205-214).
template <ELEMENT_TYPE ET>
int ElmatForm(int fe_order, int dord, const int * vnums, const double * geom, const double * lset, int dt, int intorder,
              const double * form8, double * elmat_out, LocalHeap & lh)
{
  constexpr int D = ET_trait<ET>::DIM;
  H1HighOrderFE<ET> fel(fe_order);
  if (vnums) for (int i = 0; i < D+1; i++) fel.SetVertexNumber(i, vnums[i]);
  int nd = fel.GetNDof();
  auto trafo = MakeTrafo<ET>(dord, geom, lh);
  Vector<> lv(D+1);
  for (int i = 0; i < D+1; i++) lv(i) = lset[i];
  const IntegrationRule * ir = StraightCutIntegrationRule(lv, *trafo, DOMAIN_TYPE(dt), intorder, FIND_OPTIMAL, lh);
  if (!ir) return -1;
  int np = ir->Size();
  MappedIntegrationRule<D,D> mir(*ir, *trafo, lh);
  FlatMatrix<> elmat(nd, nd, lh); elmat = 0.0;
  FlatVector<> shape(nd, lh), dn(nd, lh);
  FlatMatrixFixWidth<D> dshape(nd, lh);
  for (int q = 0; q < np; q++)
    {
      double fac = mir[q].GetMeasure() * (*ir)[q].Weight();
      double h = MeshSize<D>(mir[q]);
      fel.CalcShape((*ir)[q], shape);
      fel.CalcMappedDShape(mir[q], dshape);
      Vec<D> n = LsetNormal<ET>(mir[q], lset, lh);
      dn = dshape * n;
      double cm = form8[0] * pow(h, form8[4]) * fac, cs = form8[1] * pow(h, form8[5]) * fac;
      double cn = form8[2] * pow(h, form8[6]) * fac, cd = form8[3] * pow(h, form8[7]) * fac;
      for (int i = 0; i < nd; i++) for (int j = 0; j < nd; j++)
        {
          double gg = 0; for (int d = 0; d < D; d++) gg += dshape(i,d) * dshape(j,d);
          elmat(i,j) += cs * gg + cm * shape(i) * shape(j) + cn * (dn(j) * shape(i) + dn(i) * shape(j)) + cd * dn(i) * dn(j);
        }
    }
  for (int i = 0; i < nd; i++) for (int j = 0; j < nd; j++) elmat_out[i*nd+j] = elmat(i,j);
  return np;
}

// SymbolicFacetPatchBilinearFormIntegrator::CalcFacetMatrix (xfem/symboliccutbfi.cpp:1202-1411, time_order == -1) for the
// ghost-penalty form coef * h^hpow * (u - u.Other()) * (v - v.Other()) (py_tutorials/fictdom_ghostpen.py:105-107):
// volume rules of order 2*maxorder on both elements (:1219-1220), every point pulled into the other element by the
// reference's MapPatchIntegrationPoint (:1229-1244), weights mir1.GetWeight() (:1365), the coefficient evaluated on mir1 (:1359).
// elmat (2 nd)x(2 nd) row-major over [dofs of element 1 | dofs of element 2].  Returns the number of patch points.
template <ELEMENT_TYPE ET>
int FacetPatch(int fe_order, int dord, const int * vn1, const double * g1, const int * vn2, const double * g2,
               double coef, int hpow, double * elmat_out, double * ip2_out, LocalHeap & lh)
{
  constexpr int D = ET_trait<ET>::DIM;
  H1HighOrderFE<ET> fel1(fe_order), fel2(fe_order);
  for (int i = 0; i < D+1; i++) { if (vn1) fel1.SetVertexNumber(i, vn1[i]); if (vn2) fel2.SetVertexNumber(i, vn2[i]); }
  int nd = fel1.GetNDof(), n2 = 2*nd;
  auto trafo1 = MakeTrafo<ET>(dord, g1, lh);
  auto trafo2 = MakeTrafo<ET>(dord, g2, lh);
  IntegrationRule ir_vol1(ET, 2*fe_order);
  IntegrationRule ir_vol2(ET, 2*fe_order);
  IntegrationRule ir_patch1 (ir_vol1.Size()+ir_vol2.Size(),lh);
  IntegrationRule ir_patch2 (ir_vol1.Size()+ir_vol2.Size(),lh);
  for (int l = 0; l < ir_patch1.Size(); l++)
    {
      if (l < ir_vol1.Size())
        {
          ir_patch1[l] = ir_vol1[l];
          refpatch::MapPatchIntegrationPoint<D>(ir_patch1[l], *trafo1, *trafo2, ir_patch2[l], lh);
        }
      else
        {
          ir_patch2[l] = ir_vol2[l - ir_vol1.Size()];
          refpatch::MapPatchIntegrationPoint<D>(ir_patch2[l], *trafo2, *trafo1, ir_patch1[l], lh);
        }
    }
  MappedIntegrationRule<D,D> mir1(ir_patch1, *trafo1, lh);
  FlatMatrix<> elmat(n2, n2, lh); elmat = 0.0;
  FlatVector<> sh1(nd, lh), sh2(nd, lh), jump(n2, lh);
  for (int i = 0; i < mir1.Size(); i++)
    {
      double val = coef * pow(MeshSize<D>(mir1[i]), hpow) * mir1[i].GetWeight();
      fel1.CalcShape(ir_patch1[i], sh1);
      fel2.CalcShape(ir_patch2[i], sh2);
      jump.Range(0, nd) = sh1;
      jump.Range(nd, n2) = -sh2;
      elmat += val * jump * Trans(jump);
      if (ip2_out) for (int d = 0; d < 3; d++) { ip2_out[6*i+d] = ir_patch1[i](d); ip2_out[6*i+3+d] = ir_patch2[i](d); }
    }
  for (int i = 0; i < n2; i++) for (int j = 0; j < n2; j++) elmat_out[i*n2+j] = elmat(i,j);
  return mir1.Size();
}

// B-vectors of the compound element [H1 element | XFiniteElement] at one mapped point: values U[k] and mapped gradients G[k]
// of u_std + op_k(u_x), k = NEG, POS -- the x-part through the reference's DiffOpX (xfem/xfemdiffops.cpp:8-82)
template <ELEMENT_TYPE ET>
struct XB
{
  enum { D = ET_trait<ET>::DIM };
  T_DifferentialOperator<DiffOpX<D,DIFFOPX::RNEG>> rneg;
  T_DifferentialOperator<DiffOpX<D,DIFFOPX::RPOS>> rpos;
  T_DifferentialOperator<DiffOpX<D,DIFFOPX::RNEG_GRAD>> rneg_grad;
  T_DifferentialOperator<DiffOpX<D,DIFFOPX::RPOS_GRAD>> rpos_grad;

  void Eval(const H1HighOrderFE<ET> & fel, const XFiniteElement & xfe, const MappedIntegrationPoint<D,D> & mip,
            FlatMatrix<> U /* 2 x 2nd */, FlatMatrix<> G /* 2*D x 2nd */, LocalHeap & lh) const
  {
    HeapReset hr(lh);
    int nd = fel.GetNDof();
    FlatVector<> shape(nd, lh);
    FlatMatrixFixWidth<D> dshape(nd, lh);
    fel.CalcShape(mip.IP(), shape);
    fel.CalcMappedDShape(mip, dshape);
    FlatMatrix<double,ColMajor> bv(1, nd, lh), bg(D, nd, lh);
    for (int k = 0; k < 2; k++)
      {
        const DifferentialOperator & ov = k == 0 ? (const DifferentialOperator&)rneg : (const DifferentialOperator&)rpos;
        const DifferentialOperator & og = k == 0 ? (const DifferentialOperator&)rneg_grad : (const DifferentialOperator&)rpos_grad;
        ov.CalcMatrix(xfe, mip, bv, lh);
        og.CalcMatrix(xfe, mip, bg, lh);
        for (int i = 0; i < nd; i++)
          {
            U(k, i) = shape(i); U(k, nd+i) = bv(0, i);
            for (int d = 0; d < D; d++) { G(k*D+d, i) = dshape(i,d); G(k*D+d, nd+i) = bg(d, i); }
          }
      }
  }
};

// two-phase Nitsche-XFEM element matrix of py_tutorials/nxfem.py:77-107 on the compound element (the three integrators
// summed): NEG alpha0 gradu0.gradv0 (order 2p-2), POS alpha1 gradu1.gradv1, IF avgflux_u [v] + avgflux_v [u] + stab [u][v]
// (order 2p), stab = stab_factor (alpha0+alpha1)/h, avgflux = -sum_k kappa_k alpha_k grad u_k . n, kappa = (ratio, 1-ratio).
// xsign == nullptr: uncut element (no x-dofs): the nd x nd block of the std dofs.  elmat (2nd)x(2nd) (or nd x nd) row-major.
template <ELEMENT_TYPE ET>
int NxfemElmat(int fe_order, int dord, const int * vnums, const double * pmat, const double * lset, const signed char * xsign,
               double kappa_neg, double alpha0, double alpha1, double stab_factor, double * elmat_out, LocalHeap & lh)
{
  constexpr int D = ET_trait<ET>::DIM;
  H1HighOrderFE<ET> fel(fe_order);
  if (vnums) for (int i = 0; i < D+1; i++) fel.SetVertexNumber(i, vnums[i]);
  int nd = fel.GetNDof(), n2 = 2*nd;
  Array<DOMAIN_TYPE> signs(nd);
  for (int i = 0; i < nd; i++) signs[i] = xsign ? DOMAIN_TYPE(xsign[i]) : IF;      // IF matches neither RNEG nor RPOS: zero x-part
  XFiniteElement xfe(fel, signs, lh);
  XB<ET> xb;
  auto trafo = MakeTrafo<ET>(dord, pmat, lh);
  Vector<> lv(D+1);
  for (int i = 0; i < D+1; i++) lv(i) = lset[i];
  FlatMatrix<> A(n2, n2, lh); A = 0.0;
  FlatMatrix<> U(2, n2, lh), G(2*D, n2, lh);
  double alpha[2] = {alpha0, alpha1}, kappa[2] = {kappa_neg, 1.0 - kappa_neg};
  int total = 0;
  for (int dt = 0; dt < 2; dt++)
    {
      const IntegrationRule * ir = StraightCutIntegrationRule(lv, *trafo, DOMAIN_TYPE(dt), 2*fe_order-2, FIND_OPTIMAL, lh);
      if (!ir) continue;
      MappedIntegrationRule<D,D> mir(*ir, *trafo, lh);
      total += ir->Size();
      for (int q = 0; q < ir->Size(); q++)
        {
          xb.Eval(fel, xfe, mir[q], U, G, lh);
          double fac = mir[q].GetMeasure() * (*ir)[q].Weight() * alpha[dt];
          for (int i = 0; i < n2; i++) for (int j = 0; j < n2; j++)
            for (int d = 0; d < D; d++) A(i,j) += fac * G(dt*D+d, i) * G(dt*D+d, j);
        }
    }
  const IntegrationRule * ir = StraightCutIntegrationRule(lv, *trafo, IF, 2*fe_order, FIND_OPTIMAL, lh);
  if (ir)
    {
      MappedIntegrationRule<D,D> mir(*ir, *trafo, lh);
      total += ir->Size();
      FlatVector<> flux(n2, lh), jump(n2, lh);
      for (int q = 0; q < ir->Size(); q++)
        {
          xb.Eval(fel, xfe, mir[q], U, G, lh);
          Vec<D> n = LsetNormal<ET>(mir[q], lset, lh);
          double stab = stab_factor * (alpha0 + alpha1) / MeshSize<D>(mir[q]);
          for (int i = 0; i < n2; i++)
            {
              double f = 0;
              for (int k = 0; k < 2; k++) for (int d = 0; d < D; d++) f -= kappa[k] * alpha[k] * G(k*D+d, i) * n(d);
              flux(i) = f;
              jump(i) = U(0,i) - U(1,i);
            }
          double fac = mir[q].GetMeasure() * (*ir)[q].Weight();
          for (int i = 0; i < n2; i++) for (int j = 0; j < n2; j++)
            A(i,j) += fac * (jump(i)*flux(j) + flux(i)*jump(j) + stab * jump(i)*jump(j));
        }
    }
  int no = xsign ? n2 : nd;
  for (int i = 0; i < no; i++) for (int j = 0; j < no; j++) elmat_out[i*no+j] = A(i,j);
  return total;
}

// SymbolicLFI(levelset_domain = dt, form = f (v_std + op(v_x))), op = neg (dt NEG) / pos (dt POS), py_tutorials/nxfem.py:109-110
template <ELEMENT_TYPE ET>
int NxfemElvec(int fe_order, int dord, const int * vnums, const double * pmat, const double * lset, const signed char * xsign, int dt,
               int nmono, const double * mono, double * elvec_out, LocalHeap & lh)
{
  constexpr int D = ET_trait<ET>::DIM;
  H1HighOrderFE<ET> fel(fe_order);
  if (vnums) for (int i = 0; i < D+1; i++) fel.SetVertexNumber(i, vnums[i]);
  int nd = fel.GetNDof(), n2 = 2*nd;
  Array<DOMAIN_TYPE> signs(nd);
  for (int i = 0; i < nd; i++) signs[i] = xsign ? DOMAIN_TYPE(xsign[i]) : IF;
  XFiniteElement xfe(fel, signs, lh);
  XB<ET> xb;
  auto trafo = MakeTrafo<ET>(dord, pmat, lh);
  Vector<> lv(D+1);
  for (int i = 0; i < D+1; i++) lv(i) = lset[i];
  const IntegrationRule * ir = StraightCutIntegrationRule(lv, *trafo, DOMAIN_TYPE(dt), 2*fe_order, FIND_OPTIMAL, lh);
  FlatVector<> b(n2, lh); b = 0.0;
  int np = -1;
  if (ir)
    {
      np = ir->Size();
      MappedIntegrationRule<D,D> mir(*ir, *trafo, lh);
      FlatMatrix<> U(2, n2, lh), G(2*D, n2, lh);
      for (int q = 0; q < np; q++)
        {
          xb.Eval(fel, xfe, mir[q], U, G, lh);
          double x[3] = {0,0,0};
          for (int d = 0; d < D; d++) x[d] = mir[q].GetPoint()(d);
          double f = 0;
          for (int m = 0; m < nmono; m++)
            f += mono[4*m] * pow(x[0], mono[4*m+1]) * pow(x[1], mono[4*m+2]) * pow(x[2], mono[4*m+3]);
          double fac = mir[q].GetMeasure() * (*ir)[q].Weight() * f;
          for (int i = 0; i < n2; i++) b(i) += fac * U(dt, i);
        }
    }
  int no = xsign ? n2 : nd;
  for (int i = 0; i < no; i++) elvec_out[i] = b(i);
  return np;
}

} // namespace

extern "C" {

int ref_select_rule(int et, int order, double * pts3, double * wts, int cap)
{
  try { return CopyRule(&SelectIntegrationRule(ELEMENT_TYPE(et), order), pts3, wts, cap); }
  catch (...) { return -2; }
}

// cutint/straightcutrule.cpp:663-726
int ref_cut_rule_untrafo(int et, const double * lset, int dt, int order, double * pts3, double * wts, int cap)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  try {
    int nv = NV(ELEMENT_TYPE(et));
    Vector<> lv(nv); for (int i = 0; i < nv; i++) lv(i) = lset[i];
    auto ir = StraightCutIntegrationRuleUntransformed(lv, ELEMENT_TYPE(et), DOMAIN_TYPE(dt), order, FIND_OPTIMAL, lh);
    return CopyRule(ir, pts3, wts, cap);
  } catch (...) { return -2; }
}

// cutint/straightcutrule.cpp:580-661 (IF weights carry |F^-T n|)
int ref_cut_rule(int et, const double * pmat, const double * lset, int dt, int order, double * pts3, double * wts, int cap)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  try {
    const IntegrationRule * ir = nullptr;
    if (et == ET_SEGM) { FE_ElementTransformation<1,1> * t; ir = TrafoRule<1>(ET_SEGM, pmat, lset, dt, order, lh, t); }
    else if (et == ET_TRIG) { FE_ElementTransformation<2,2> * t; ir = TrafoRule<2>(ET_TRIG, pmat, lset, dt, order, lh, t); }
    else if (et == ET_TET) { FE_ElementTransformation<3,3> * t; ir = TrafoRule<3>(ET_TET, pmat, lset, dt, order, lh, t); }
    else return -4;
    return CopyRule(ir, pts3, wts, cap);
  } catch (...) { return -2; }
}

// deformed element: StraightCutIntegrationRule on an isoparametric trafo (IF weights: per-point |F^-T n|,
// cutint/straightcutrule.cpp:558-576), measure and mapped point of every rule point
int ref_cut_rule_iso(int et, int dord, const double * gcoef, const double * lset, int dt, int order,
                     double * pts3, double * wts, double * meas, double * xphys, int cap)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  try {
    if (et == ET_TRIG) return RuleIso<ET_TRIG>(dord, gcoef, lset, dt, order, pts3, wts, meas, xphys, cap, lh);
    if (et == ET_TET) return RuleIso<ET_TET>(dord, gcoef, lset, dt, order, pts3, wts, meas, xphys, cap, lh);
    return -4;
  } catch (...) { return -2; }
}

int ref_elmat_h1_iso(int et, int fe_order, int dord, const double * gcoef, const double * lset, int dt, int intorder,
                     double c_mass, double c_stiff, double * elmat)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  try {
    if (et == ET_TRIG) return ElmatIso<ET_TRIG>(fe_order, dord, gcoef, lset, dt, intorder, c_mass, c_stiff, elmat, lh);
    if (et == ET_TET) return ElmatIso<ET_TET>(fe_order, dord, gcoef, lset, dt, intorder, c_mass, c_stiff, elmat, lh);
    return -4;
  } catch (...) { return -2; }
}

int ref_elvec_h1_iso(int et, int fe_order, int dord, const double * gcoef, const double * lset, int dt, int intorder,
                     int nmono, const double * mono, double * elvec)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  try {
    if (et == ET_TRIG) return ElvecIso<ET_TRIG>(fe_order, dord, gcoef, lset, dt, intorder, nmono, mono, elvec, lh);
    if (et == ET_TET) return ElvecIso<ET_TET>(fe_order, dord, gcoef, lset, dt, intorder, nmono, mono, elvec, lh);
    return -4;
  } catch (...) { return -2; }
}

// cutint/spacetimecutrule.cpp:98-196 with the nodal time element of order ntdofs - 1 (spacetime/SpaceTimeFE.cpp:116-173).
// lset_st [ntdofs][nv]: time-major like SpaceTimeFESpace::GetDofNrs (spacetime/SpaceTimeFESpace.cpp:109-122).
// out: spatial reference points, reference time of each point (the reference smuggles it through Weight()), weights (wei_arr)
int ref_st_cut_rule(int et, const double * pmat, const double * lset_st, int ntdofs, int dt, int order_time, int order_space,
                    double * pts3, double * tq, double * wts, int cap)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  try {
    if (et != ET_TRIG && et != ET_TET) return -4;
    const int D = et == ET_TRIG ? 2 : 3, nv = D + 1;
    Vector<> lv(ntdofs * nv);
    for (int i = 0; i < ntdofs * nv; i++) lv(i) = lset_st[i];
    NodalTimeFE tfe(ntdofs - 1, false, false);
    const IntegrationRule * ir = nullptr;
    Array<double> wei;
    Matrix<> pm(nv, D);
    for (int i = 0; i < nv; i++) for (int d = 0; d < D; d++) pm(i,d) = pmat[i*D+d];
    if (D == 2) {
      auto trafo = new (lh) FE_ElementTransformation<2,2>(ET_TRIG, pm);
      tie(ir, wei) = SpaceTimeCutIntegrationRule(lv, *trafo, &tfe, DOMAIN_TYPE(dt), order_time, order_space, FIND_OPTIMAL, lh);
    } else {
      auto trafo = new (lh) FE_ElementTransformation<3,3>(ET_TET, pm);
      tie(ir, wei) = SpaceTimeCutIntegrationRule(lv, *trafo, &tfe, DOMAIN_TYPE(dt), order_time, order_space, FIND_OPTIMAL, lh);
    }
    if (!ir) return -1;
    const int n = ir->Size();
    if (n > cap) return -3;
    for (int i = 0; i < n; i++) {
      for (int d = 0; d < 3; d++) pts3[3*i+d] = (*ir)[i](d);
      tq[i] = (*ir)[i].Weight();
      wts[i] = wei[i];
    }
    return n;
  } catch (...) { return -2; }
}

// space-time element matrix through the reference's own SpaceTimeFE (spacetime/SpaceTimeFE.cpp:29-113: CalcShape, CalcDShape,
// CalcDtShape on the integration points of SpaceTimeCutIntegrationRule, which carry the reference time in Weight()) over
// H1HighOrderFE x NodalTimeFE; the block loop  elmat += |det J| w (c_m u v + c_s grad u . grad v + c_t dt(u) v)  is this driver's.
}  // extern "C"
template <int D, ELEMENT_TYPE ET>
static int st_elmat_t(int fe_order, int ntd_fe, const double * pmat, const Vector<> & lv, int ntdofs, int dt, int order_time, int order_space,
                      double cm, double cs, double ct, double * elmat, LocalHeap & lh)
{
  const int nv = D + 1;
  NodalTimeFE tfe_lset(ntdofs - 1, false, false);
  NodalTimeFE tfe(ntd_fe - 1, false, false);
  H1HighOrderFE<ET> sfe(fe_order);
  SpaceTimeFE<D> stfe(&sfe, &tfe, false, 0.0);
  Matrix<> pm(nv, D);
  for (int i = 0; i < nv; i++) for (int d = 0; d < D; d++) pm(i,d) = pmat[i*D+d];
  auto trafo = new (lh) FE_ElementTransformation<D,D>(ET, pm);
  const IntegrationRule * ir = nullptr;
  Array<double> wei;
  tie(ir, wei) = SpaceTimeCutIntegrationRule(lv, *trafo, &tfe_lset, DOMAIN_TYPE(dt), order_time, order_space, FIND_OPTIMAL, lh);
  const int N = stfe.GetNDof();
  for (int i = 0; i < N*N; i++) elmat[i] = 0.0;
  if (!ir) return -1;
  Vector<> shape(N), dtshape(N);
  Matrix<> dshape(N, D), dphys(N, D);
  for (int q = 0; q < ir->Size(); q++) {
    const IntegrationPoint & ip = (*ir)[q];
    MappedIntegrationPoint<D,D> mip(ip, *trafo);
    stfe.CalcShape(ip, shape);
    stfe.CalcDShape(ip, dshape);
    stfe.CalcDtShape(ip, dtshape);
    dphys = dshape * mip.GetJacobianInverse();
    const double fac = fabs(mip.GetJacobiDet()) * wei[q];
    for (int i = 0; i < N; i++)
      for (int j = 0; j < N; j++) {
        double g = 0.0;
        for (int d = 0; d < D; d++) g += dphys(i,d) * dphys(j,d);
        elmat[i*N+j] += fac * (cm * shape(i) * shape(j) + cs * g + ct * shape(i) * dtshape(j));
      }
  }
  return ir->Size();
}

extern "C" {
int ref_st_elmat_h1(int et, int fe_order, int ntd_fe, const double * pmat, const double * lset_st, int ntdofs, int dt, int order_time,
                    int order_space, double cm, double cs, double ct, double * elmat)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  try {
    if (et != ET_TRIG && et != ET_TET) return -4;
    const int nv = et == ET_TRIG ? 3 : 4;
    Vector<> lv(ntdofs * nv);
    for (int i = 0; i < ntdofs * nv; i++) lv(i) = lset_st[i];
    if (et == ET_TRIG) return st_elmat_t<2, ET_TRIG>(fe_order, ntd_fe, pmat, lv, ntdofs, dt, order_time, order_space, cm, cs, ct, elmat, lh);
    return st_elmat_t<3, ET_TET>(fe_order, ntd_fe, pmat, lv, ntdofs, dt, order_time, order_space, cm, cs, ct, elmat, lh);
  } catch (...) { return -2; }
}

// xfem/cutinfo.cpp:79-111 : two order-0 rules -> part volumes -> domain + cut ratio.  returns DOMAIN_TYPE or -2
int ref_cutinfo_element(int et, const double * pmat, const double * lset, double * ratio)
{
  double part_vol[2] = {0.0, 0.0};
  double pts[3*64], wts[64];
  for (int np : {1, 0})
    {
      int n = ref_cut_rule(et, pmat, lset, np, 0, pts, wts, 64);
      if (n == -2) return -2;
      for (int i = 0; i < n; i++) part_vol[np] += wts[i];
    }
  *ratio = part_vol[0] / (part_vol[0] + part_vol[1]);
  if (part_vol[0] > 0.0) return part_vol[1] > 0.0 ? 2 : 0;
  return 1;
}

int ref_h1_ndof(int et, int order)
{
  if (et == ET_TRIG) return H1HighOrderFE<ET_TRIG>(order).GetNDof();
  if (et == ET_TET) return H1HighOrderFE<ET_TET>(order).GetNDof();
  if (et == ET_SEGM) return H1HighOrderFE<ET_SEGM>(order).GetNDof();
  return -1;
}

// shape (ndof) and reference-gradient dshape (ndof x dim) at a reference point
int ref_h1_shape(int et, int order, const int * vnums, const double * ip3, double * shape, double * dshape)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  IntegrationPoint ip(ip3[0], ip3[1], ip3[2], 1.0);
  if (et == ET_TRIG)
    {
      H1HighOrderFE<ET_TRIG> fel(order);
      if (vnums) for (int i = 0; i < 3; i++) fel.SetVertexNumber(i, vnums[i]);
      int nd = fel.GetNDof();
      FlatVector<> s(nd, lh); FlatMatrixFixWidth<2> ds(nd, lh);
      fel.CalcShape(ip, s); fel.CalcDShape(ip, ds);
      for (int i = 0; i < nd; i++) { shape[i] = s(i); for (int d = 0; d < 2; d++) dshape[2*i+d] = ds(i,d); }
      return nd;
    }
  if (et == ET_TET)
    {
      H1HighOrderFE<ET_TET> fel(order);
      if (vnums) for (int i = 0; i < 4; i++) fel.SetVertexNumber(i, vnums[i]);
      int nd = fel.GetNDof();
      FlatVector<> s(nd, lh); FlatMatrixFixWidth<3> ds(nd, lh);
      fel.CalcShape(ip, s); fel.CalcDShape(ip, ds);
      for (int i = 0; i < nd; i++) { shape[i] = s(i); for (int d = 0; d < 3; d++) dshape[3*i+d] = ds(i,d); }
      return nd;
    }
  return -1;
}

int ref_elmat_h1(int et, int fe_order, const int * vnums, const double * pmat, const double * lset, int dt, int intorder,
                 double c_mass, double c_stiff, double * elmat)
{
  LocalHeap & lh = GetLH();
  try {
    if (et == ET_TRIG) return ElmatWork<ET_TRIG>::Elmat(fe_order, vnums, pmat, lset, dt, intorder, c_mass, c_stiff, elmat, lh);
    if (et == ET_TET) return ElmatWork<ET_TET>::Elmat(fe_order, vnums, pmat, lset, dt, intorder, c_mass, c_stiff, elmat, lh);
    return -4;
  } catch (...) { return -2; }
}

int ref_elvec_h1(int et, int fe_order, const int * vnums, const double * pmat, const double * lset, int dt, int intorder,
                 int nmono, const double * mono, double * elvec)
{
  LocalHeap & lh = GetLH();
  try {
    if (et == ET_TRIG) return ElmatWork<ET_TRIG>::Elvec(fe_order, vnums, pmat, lset, dt, intorder, nmono, mono, elvec, lh);
    if (et == ET_TET) return ElmatWork<ET_TET>::Elvec(fe_order, vnums, pmat, lset, dt, intorder, nmono, mono, elvec, lh);
    return -4;
  } catch (...) { return -2; }
}

// ---------------------------------------------------------------------------------------------------------------
// batched mesh-level drivers (CPU baseline; OpenMP threads stand in for netgen's TaskManager workers, each with its
// own LocalHeap exactly like IterateRange's lh.Split, utils/ngsxstd.cpp:3-31)
// ---------------------------------------------------------------------------------------------------------------

static void GatherElement(int dim, const int * el2v, const double * vert, const double * lset_v, long e, double * pmat, double * lset)
{
  for (int i = 0; i < dim+1; i++)
    {
      int v = el2v[e*(dim+1)+i];
      for (int d = 0; d < dim; d++) pmat[i*dim+d] = vert[(long)v*dim+d];
      lset[i] = lset_v[v];
    }
}

// CutInformation::Update element loop.  dom[e] in {0,1,2}; returns #elements on which the reference threw
long ref_cutinfo_mesh(int dim, long ne, const int * el2v, const double * vert, const double * lset_v,
                      signed char * dom, double * ratio, int nthreads)
{
  int et = dim == 2 ? ET_TRIG : ET_TET;
  long nerr = 0;
  if (nthreads > 0) omp_set_num_threads(nthreads);
#pragma omp parallel for schedule(static) reduction(+:nerr)
  for (long e = 0; e < ne; e++)
    {
      double pmat[12], lset[4];
      GatherElement(dim, el2v, vert, lset_v, e, pmat, lset);
      int r = ref_cutinfo_element(et, pmat, lset, &ratio[e]);
      if (r < 0) { nerr++; dom[e] = -1; } else dom[e] = r;
    }
  return nerr;
}

// IntegrateX (cutint/python_cutint.cpp:41-68) with cf = sum_m coef_m x^a y^b z^c; deterministic per-element partials summed serially
double ref_integrate_mesh(int dim, long ne, const int * el2v, const double * vert, const double * lset_v, int dt, int order,
                          int nmono, const double * mono, int nthreads, long * npts_total)
{
  int et = dim == 2 ? ET_TRIG : ET_TET;
  if (nthreads > 0) omp_set_num_threads(nthreads);
  long double sum = 0; long np = 0;      // extended accumulator: the comparison should see the integrand, not 10^7 roundings of the checker
#pragma omp parallel for schedule(static) reduction(+:sum,np)
  for (long e = 0; e < ne; e++)
    {
      LocalHeap & lh = GetLH(); HeapReset hr(lh);
      double pmat[12], lset[4];
      GatherElement(dim, el2v, vert, lset_v, e, pmat, lset);
      double lsum = 0; int n = 0;
      auto work = [&](auto DD)
        {
          constexpr int D = decltype(DD)::value;
          FE_ElementTransformation<D,D> * trafo;
          const IntegrationRule * ir = TrafoRule<D>(ELEMENT_TYPE(et), pmat, lset, dt, order, lh, trafo);
          if (!ir) return;
          n = ir->Size();
          MappedIntegrationRule<D,D> mir(*ir, *trafo, lh);
          for (int i = 0; i < n; i++)
            {
              double x[3] = {0,0,0};
              for (int d = 0; d < D; d++) x[d] = mir[i].GetPoint()(d);
              double f = 0;
              for (int m = 0; m < nmono; m++)
                f += mono[4*m] * pow(x[0], mono[4*m+1]) * pow(x[1], mono[4*m+2]) * pow(x[2], mono[4*m+3]);
              lsum += mir[i].GetMeasure() * (*ir)[i].Weight() * f;
            }
        };
      if (dim == 2) work(std::integral_constant<int,2>()); else work(std::integral_constant<int,3>());
      sum += lsum; np += n;
    }
  if (npts_total) *npts_total = np;
  return (double)sum;
}

// element matrices for a list of elements (els == nullptr: all), written densely elmats[k*nd*nd]; npts[k] = rule size or -1
long ref_elmat_mesh(int dim, long nel, const long * els, const int * el2v, const double * vert, const double * lset_v,
                    int fe_order, int dt, int intorder, double c_mass, double c_stiff, double * elmats, int * npts, int nthreads)
{
  int et = dim == 2 ? ET_TRIG : ET_TET;
  int nd = ref_h1_ndof(et, fe_order);
  if (nthreads > 0) omp_set_num_threads(nthreads);
  long nnz = 0;
#pragma omp parallel for schedule(dynamic,256) reduction(+:nnz)
  for (long k = 0; k < nel; k++)
    {
      long e = els ? els[k] : k;
      double pmat[12], lset[4];
      GatherElement(dim, el2v, vert, lset_v, e, pmat, lset);
      int vn[4]; for (int i = 0; i < dim+1; i++) vn[i] = el2v[e*(dim+1)+i];
      int r = ref_elmat_h1(et, fe_order, vn, pmat, lset, dt, intorder, c_mass, c_stiff, elmats + (size_t)k*nd*nd);
      if (npts) npts[k] = r;
      if (r < 0) for (int i = 0; i < nd*nd; i++) elmats[(size_t)k*nd*nd+i] = 0.0;
      else nnz++;
    }
  return nnz;
}

// Full assemble-to-CSR on the host: reference element matrix per element, then the row-search "+=" of
// SparseMatrixTM::AddElementMatrix (NGS/linalg/sparsematrix_impl.hpp:63-103: columns sorted, linear scan of the CSR row),
// colour by colour (NGS/comp/fespace.cpp:1540-1566: one parallel loop per colour).
long ref_assemble_csr(int dim, const int * el2v, const double * vert, const double * lset_v,
                      int fe_order, const int * el2dof, int dt, int intorder, double c_mass, double c_stiff,
                      int ncolors, const long * color_ptr, const int * color_els,
                      const long * rowptr, const int * colidx, double * vals, int nthreads)
{
  int et = dim == 2 ? ET_TRIG : ET_TET;
  int nd = ref_h1_ndof(et, fe_order);
  if (nthreads > 0) omp_set_num_threads(nthreads);
  long nass = 0;
  for (int c = 0; c < ncolors; c++)
    {
#pragma omp parallel for schedule(dynamic,64) reduction(+:nass)
      for (long k = color_ptr[c]; k < color_ptr[c+1]; k++)
        {
          long e = color_els[k];
          double pmat[12], lset[4], elmat[400];
          GatherElement(dim, el2v, vert, lset_v, e, pmat, lset);
          int vn[4]; for (int i = 0; i < dim+1; i++) vn[i] = el2v[e*(dim+1)+i];
          int r = ref_elmat_h1(et, fe_order, vn, pmat, lset, dt, intorder, c_mass, c_stiff, elmat);
          if (r < 0) continue;
          nass++;
          const int * dn = el2dof + e*nd;
          int map[20]; for (int i = 0; i < nd; i++) map[i] = i;
          std::sort(map, map+nd, [&](int a, int b) { return dn[a] < dn[b]; });
          for (int i = 0; i < nd; i++)
            {
              long k0 = rowptr[dn[i]];
              for (int j1 = 0; j1 < nd; j1++)
                {
                  int j = map[j1];
                  while (colidx[k0] != dn[j]) k0++;
                  vals[k0] += elmat[i*nd+j];
                }
            }
        }
    }
  return nass;
}

// general ngsx_form_h1 element matrix (Nitsche / normal-normal terms, powers of mesh_size) on an affine (dord = 0) or
// isoparametric element
int ref_elmat_h1_form(int et, int fe_order, int dord, const int * vnums, const double * geom, const double * lset, int dt, int intorder,
                      const double * form8, double * elmat)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  try {
    if (et == ET_TRIG) return ElmatForm<ET_TRIG>(fe_order, dord, vnums, geom, lset, dt, intorder, form8, elmat, lh);
    if (et == ET_TET) return ElmatForm<ET_TET>(fe_order, dord, vnums, geom, lset, dt, intorder, form8, elmat, lh);
    return -4;
  } catch (...) { return -2; }
}

// facet-patch (ghost penalty) matrix; ip2 (may be NULL) [npts][6]: the patch points in the reference coordinates of both elements
int ref_facetpatch_elmat_h1(int et, int fe_order, int dord, const int * vn1, const double * g1, const int * vn2, const double * g2,
                            double coef, int hpow, double * elmat, double * ip2)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  try {
    if (et == ET_TRIG) return FacetPatch<ET_TRIG>(fe_order, dord, vn1, g1, vn2, g2, coef, hpow, elmat, ip2, lh);
    if (et == ET_TET) return FacetPatch<ET_TET>(fe_order, dord, vn1, g1, vn2, g2, coef, hpow, elmat, ip2, lh);
    return -4;
  } catch (...) { return -2; }
}

// dord == 0: affine element, geom = vertex matrix; dord >= 1: isoparametric element of that order, geom = geometry coefficients
int ref_nxfem_elmat(int et, int fe_order, int dord, const int * vnums, const double * pmat, const double * lset, const signed char * xsign,
                    double kappa_neg, double alpha0, double alpha1, double stab_factor, double * elmat)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  try {
    if (et == ET_TRIG) return NxfemElmat<ET_TRIG>(fe_order, dord, vnums, pmat, lset, xsign, kappa_neg, alpha0, alpha1, stab_factor, elmat, lh);
    if (et == ET_TET) return NxfemElmat<ET_TET>(fe_order, dord, vnums, pmat, lset, xsign, kappa_neg, alpha0, alpha1, stab_factor, elmat, lh);
    return -4;
  } catch (...) { return -2; }
}

int ref_nxfem_elvec(int et, int fe_order, int dord, const int * vnums, const double * pmat, const double * lset, const signed char * xsign, int dt,
                    int nmono, const double * mono, double * elvec)
{
  LocalHeap & lh = GetLH(); HeapReset hr(lh);
  try {
    if (et == ET_TRIG) return NxfemElvec<ET_TRIG>(fe_order, dord, vnums, pmat, lset, xsign, dt, nmono, mono, elvec, lh);
    if (et == ET_TET) return NxfemElvec<ET_TET>(fe_order, dord, vnums, pmat, lset, xsign, dt, nmono, mono, elvec, lh);
    return -4;
  } catch (...) { return -2; }
}

int ref_max_threads() { return omp_get_max_threads(); }

} // extern "C"
