/* ngsx_abi.h -- C ABI of libngsx_b200.so: the B200-native (sm_100a) cut-FEM assembly hot path of ngsxfem.
 *
 * Batched replacement for the per-element virtual calls of the reference: one call per
 * CutInfo.Update / Assemble / Integrate over flat device arrays.  Plain pointers and sizes only; no C++/torch
 * types cross this boundary.  All pointers are DEVICE pointers unless the name starts with h_.  Every entry point
 * returns 0 on success or a negative ngsx_status; ngsx_last_error() returns a thread-local message.  `stream` is a
 * cudaStream_t passed as void* (NULL = legacy default stream); calls are asynchronous on it unless noted.
 *
 * Reference interfaces replaced (paths relative to the ngsxfem tree; NGS/ = external_dependencies/ngsolve/):
 *   enums                     utils/ngsxstd.hpp:8-78, python/python_ngsxfem.cpp:14-37
 *   ngsx_interpolate_p1_*     utils/p1interpol.cpp:25-96            (InterpolateToP1, utils/python_utils.cpp:21-81)
 *   ngsx_classify             xfem/cutinfo.cpp:54-120               (CutInfo.Update, xfem/python_xfem.cpp:57-70)
 *   ngsx_node_marks           xfem/cutinfo.cpp:122-186
 *   ngsx_rule_count/_fill     cutint/xintegration.cpp:13-80, cutint/straightcutrule.cpp:580-726
 *   ngsx_elmat_h1             xfem/symboliccutbfi.cpp:73-315        (CalcElementMatrixAdd, xfem/symboliccutbfi.hpp:42-58)
 *   ngsx_elvec_h1             xfem/symboliccutlfi.cpp:52-158        (CalcElementVector, xfem/symboliccutlfi.hpp:38-55)
 *   ngsx_csr_graph_*          utils/restrictedblf.cpp:22-138        (RestrictedBilinearForm::GetGraph)
 *   ngsx_color_groups         NGS/comp/fespace.cpp:598-705          (FESpace element colouring)
 *   ngsx_assemble_csr_h1      NGS/comp/bilinearform.cpp:1117-1575 + NGS/linalg/sparsematrix_impl.hpp:63-103
 *   ngsx_assemble_vec_h1      NGS/comp/linearform.cpp:209-243
 *   ngsx_integrate            cutint/python_cutint.cpp:20-72        (IntegrateX)
 *   ngsx_st_rule_count/_fill, ngsx_st_integrate, ngsx_st_elmat_h1, ngsx_st_classify
 *                             cutint/spacetimecutrule.cpp:6-196, cutint/xintegration.cpp:38-58 (time_order >= 0)
 *   ngsx_facets_neighbor_types[_periodic], ngsx_elements_with_neighbor_facets, ngsx_dofs_of_elements
 *                             xfem/cutinfo.cpp:191-329              (xfem/python_xfem.cpp:168-318; GetDofsOfFacets = ngsx_dofs_of_elements
 *                                                                    on the facet -> dof table)
 *   ngsx_facetpatch_elmat_h1, ngsx_facetpatch_cmat_h1
 *                             xfem/symboliccutbfi.cpp:1115-1411     (CalcFacetMatrix, xfem/symboliccutbfi.hpp:253-259)
 *   ngsx_xfem_elmat_h1        xfem/xfemdiffops.cpp:8-66 + xfem/symboliccutbfi.cpp:73-315 (compound [Vh, XFESpace] element)
 *   ngsx_facetpatch_table_h1, ngsx_assemble_csr_rows_mats
 *                             NGS/comp/bilinearform.cpp:1800-1853 (facet loop: compressed dofs + AddElementMatrix)
 *
 * Bit arrays are uint32 words, bit i of the array = word i/32, bit i%32: layout-compatible with netgen's
 * ngcore::BitArray (NG/libsrc/core/bitarray.hpp:150-155: byte i/8, mask 1<<(i%8)) on little-endian hosts.
 */
#ifndef NGSX_ABI_H
#define NGSX_ABI_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NGSX_ABI_VERSION 1

/* utils/ngsxstd.hpp:14-17 */
enum ngsx_domain_type { NGSX_NEG = 0, NGSX_POS = 1, NGSX_IF = 2 };
/* utils/ngsxstd.hpp:24-33 (bit mask: NEG=1, POS=2, IF=4) */
enum ngsx_combined_domain_type {
  NGSX_CDOM_NO = 0, NGSX_CDOM_NEG = 1, NGSX_CDOM_POS = 2, NGSX_CDOM_UNCUT = 3,
  NGSX_CDOM_IF = 4, NGSX_CDOM_HASNEG = 5, NGSX_CDOM_HASPOS = 6, NGSX_CDOM_ANY = 7 };

enum ngsx_status {
  NGSX_OK = 0,
  NGSX_ERR_ARG = -1,          /* bad argument (dimension, order, null pointer, ...) */
  NGSX_ERR_UNSUPPORTED = -2,  /* form / element / order outside the CUDA path: no CPU fallback by charter */
  NGSX_ERR_CUDA = -3,         /* CUDA runtime error (message has cudaGetErrorString) */
  NGSX_ERR_DEGENERATE = -4,   /* level set the reference throws on (vertex value exactly 0 on a cut element, all-zero element) */
  NGSX_ERR_CAPACITY = -5      /* caller buffer too small / too many colours */
};

typedef struct ngsx_mesh {
  int32_t dim;              /* 2: triangles (ET_TRIG), 3: tetrahedra (ET_TET) */
  int32_t _pad;
  int64_t ne, nv;
  const double  *vert;      /* [nv][dim] */
  const int32_t *el2v;      /* [ne][dim+1], local vertex order of the reference mesh */
} ngsx_mesh;

/* bilinear form on scalar H1:  sum of
 *   c_stiff * h^hpow_stiff * grad(u).grad(v) + c_mass * h^hpow_mass * u v
 * + c_nitsche * h^hpow_nitsche * ( (grad(u).n) v + (grad(v).n) u ),   n = grad(lset_p1)/|grad(lset_p1)|,  h = |det J|^(1/dim)
 * + c_dndn * h^hpow_dndn * (grad(u).n)(grad(v).n)      (c_stiff = -c_dndn: the tangential stiffness P grad(u).P grad(v),
 *                                                       P = I - n n^T, of py_tutorials/tracefem3d.py:74-88)
 * (the coefficient functions of py_tutorials/fictdom_ghostpen.py:86-96; elementwise constant => evaluated once per element
 * like xfem/symboliccutbfi.cpp:205-214). */
typedef struct ngsx_form_h1 {
  double  c_mass, c_stiff, c_nitsche, c_dndn;
  int32_t hpow_mass, hpow_stiff, hpow_nitsche, hpow_dndn;
} ngsx_form_h1;

/* polynomial coefficient function  f(x) = sum_{m<nterms} coef[m] * x^ex[m][0] * y^ex[m][1] * z^ex[m][2] */
#define NGSX_POLY_MAXTERMS 8
typedef struct ngsx_poly {
  int32_t nterms, _pad;
  double  coef[NGSX_POLY_MAXTERMS];
  int32_t ex[NGSX_POLY_MAXTERMS][3];
} ngsx_poly;

/* analytic level sets for the CoefficientFunction branch of InterpolateToP1 */
enum ngsx_lset_kind {
  NGSX_LSET_AFFINE = 0,   /* p[0] + p[1] x + p[2] y + p[3] z */
  NGSX_LSET_SPHERE = 1,   /* |x - (p[0],p[1],p[2])| - p[3]            (2D: circle, p[2] ignored) */
  NGSX_LSET_RING   = 2,   /* | |x - c| - p[3] | - p[4]                (fictdom_ghostpen.py:28-33: rc, rr) */
  NGSX_LSET_CYLZ   = 3    /* sqrt((x-p0)^2+(y-p1)^2) - p[3]           (tests/pytests/test_xfes_ndof.py:16) */
};

int         ngsx_abi_version(void);
const char *ngsx_last_error(void);
/* number of kernels launched by this library in the calling process since load (bench.py "gpu_launches") */
int64_t     ngsx_launch_count(void);

/* ---- (a-0) InterpolateToP1 ------------------------------------------------------------------------------------ */
/* GridFunction branch: out[v] = in[v], |out| < eps -> eps   (utils/p1interpol.cpp:74-90) */
int ngsx_interpolate_p1_clamp(int64_t nv, const double *in, double eps, double *out, void *stream);
/* CoefficientFunction branch for an analytic level set evaluated at the vertices, then the same clamp */
int ngsx_interpolate_p1_cf(const ngsx_mesh *mesh, int kind, const double *h_params /*[8]*/, double eps, double *out, void *stream);

/* ---- (a) CutInfo.Update ---------------------------------------------------------------------------------------- */
/* cdom_words: [8][nwords] with nwords = (ne+31)/32, indexed by ngsx_combined_domain_type (row NO stays 0, ANY all ones);
 * dom: [ne] int8 ngsx_domain_type; ratio: [ne] cut ratio vol_NEG/(vol_NEG+vol_POS) of xfem/cutinfo.cpp:91;
 * err: [2] int64 {number of degenerate elements, smallest such element number or -1}.  Any output may be NULL.
 * Degenerate elements (the reference throws) are counted in err and get dom = -1, no bits. */
int ngsx_classify(const ngsx_mesh *mesh, const double *lset_v, uint32_t *cdom_words, int8_t *dom, double *ratio,
                  int64_t *err, void *stream);
/* the BND pass of CutInformation::Update (xfem/cutinfo.cpp:47, 67: `for (VorB vb : {VOL, BND})`): boundary elements of a mesh
 * of dimension dim (segments / triangles), bnd_el2v [nbe][dim]; outputs as ngsx_classify */
int ngsx_classify_bnd(int dim, int64_t nbe, const int32_t *bnd_el2v, const double *lset_v, uint32_t *cdom_words, int8_t *dom,
                      double *ratio, int64_t *err, void *stream);
/* cut_neighboring_node / dom_of_node for one node type (xfem/cutinfo.cpp:122-186): el2node [ne][npe];
 * cut_nb_words [(nnodes+31)/32] is OR-ed (must be zeroed by the caller), dom_of_node [nnodes] int32 must be pre-filled
 * with IF(2); it is lowered with an integer atomicMin (NEG < POS < IF), which makes the reference's racy last-writer
 * store (cutinfo.cpp:163-183) deterministic.  cell_mode != 0: node == element, dom_of_node = ratio > 0.5 ? NEG : POS (:176-181).
 * err[1] of ngsx_classify / ngsx_rule_count is INT64_MAX when no element is degenerate. */
int ngsx_node_marks(int64_t ne, int32_t npe, const int32_t *el2node, const int8_t *dom, const double *ratio,
                    int64_t nnodes, int cell_mode, uint32_t *cut_nb_words, int32_t *dom_of_node, void *stream);

/* ---- (b) straight-cut quadrature ------------------------------------------------------------------------------- */
/* npts[e] = size of the reference's rule for (dt, order) on element e; 0 where the reference returns nullptr.
 * err as in ngsx_classify (may be NULL). */
int ngsx_rule_count(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, int32_t *npts, int64_t *err, void *stream);
/* out[i] = sum_{k<i} in[k] (n+1 outputs, out[n] = total) */
int ngsx_exclusive_scan(int64_t n, const int32_t *in, int64_t *out, void *stream);
/* points (reference coordinates, [total][3]) and weights ([total]) at offs[e], in the reference's emission order.
 * transform_if != 0: IF weights carry |F^-T n| like StraightCutIntegrationRule (straightcutrule.cpp:558-576);
 *               == 0: StraightCutIntegrationRuleUntransformed (:663-726).
 * emit_uncut == 0: uncut elements (whose rule is NGSolve's static table) are skipped (their offs range must be empty). */
int ngsx_rule_fill(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, int transform_if, int emit_uncut,
                   const int64_t *offs, double *pts3, double *wts, void *stream);

/* ---- (c) element matrices / vectors ---------------------------------------------------------------------------- */
/* els: [nlist] element numbers (NULL: 0..nlist-1).  elmats: [nlist][nd][nd] row-major (rows = test functions), zero where
 * the reference adds nothing; npts (may be NULL): rule size used per listed element.  fe_order 1..3, H1HighOrderFE basis (order 3:
 * vertex-oriented edge functions from the global vertex numbers in el2v + face / cell functions, NGS/fem/h1hofe_impl.hpp:78-108, 250-312;
 * ngsx_elmat_h1, ngsx_assemble_csr_h1, ngsx_elvec_h1 and ngsx_assemble_vec_h1 -- the row-gather entry point, XFEM blocks, facet patches and
 * deformed meshes stay at 1..2).
 * intorder < 0: the reference's default 2*fe_order - (difforders) is NOT guessed here -- the caller passes it. */
int ngsx_elmat_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, const ngsx_form_h1 *h_form,
                  int64_t nlist, const int32_t *els, double *elmats, int32_t *npts, void *stream);
/* the same with a GridFunction factor multiplying the whole form (the reference evaluates the coefficient tree per point,
 * xfem/symboliccutbfi.cpp:178-215): gf in H1 of order cf_order (1, 2) on the same mesh.  H1 trial / test order 1-2; every listed element
 * goes through the quadrature (no constant-rule shortcut: the coefficient varies inside the element). */
int ngsx_elmat_h1_cf(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, const ngsx_form_h1 *h_form, int cf_order,
                     const int32_t *cf_el2dof, const double *cf_vec, int64_t nlist, const int32_t *els, double *elmats, int32_t *npts, void *stream);
int ngsx_elvec_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, const ngsx_poly *h_f,
                  int64_t nlist, const int32_t *els, double *elvecs, int32_t *npts, void *stream);
/* the same with a GridFunction factor in the coefficient (the reference evaluates an arbitrary CoefficientFunction tree per point,
 * xfem/symboliccutlfi.cpp:120-140; a GridFunction CF is sum_i c_i phi_i at the point): f(x) = h_f(x) * gf(x), gf in H1 of order
 * cf_order (1, 2) on the same mesh, coefficients cf_vec [ndof_cf], dof table cf_el2dof [ne][nd_cf].  H1 test space of order 1-2. */
int ngsx_elvec_h1_cf(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, const ngsx_poly *h_f, int cf_order,
                     const int32_t *cf_el2dof, const double *cf_vec, int64_t nlist, const int32_t *els, double *elvecs, int32_t *npts, void *stream);
/* IntegrateX (cutint/python_cutint.cpp:20-72) with a GridFunction factor in the integrand: out[0] = sum_e sum_q |det J_e| w_q h_f(x_q) gf(x_q),
 * gf in H1 of order cf_order (1, 2).  Per-element shares written once each, then a reduction of fixed shape (bitwise reproducible). */
int ngsx_integrate_cf(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, const ngsx_poly *h_f, int cf_order, const int32_t *cf_el2dof,
                      const double *cf_vec, double *out, void *stream);

/* ---- (d) CSR graph, colouring, scatter ------------------------------------------------------------------------- */
typedef struct ngsx_graph ngsx_graph;   /* opaque, host object owning temporary device buffers */
int ngsx_csr_graph_create(int64_t ndof, ngsx_graph **out);
/* add the couplings of a dof table [n][nd] (element dofs of marked elements, or the merged dofs of facet patches) */
/* symmetric storage (BilinearForm(symmetric=True): SparseMatrixSymmetric keeps the lower triangle, AddElementMatrixSymmetric,
 * NGS/linalg/sparsematrix_impl.hpp:605-690): call before the first add_table; the graph then keeps the couplings col <= row only */
int ngsx_csr_graph_set_lower(ngsx_graph *g, int on);
int ngsx_csr_graph_add_table(ngsx_graph *g, int64_t n, int32_t nd, const int32_t *table, void *stream);
/* sort + unique; synchronises the stream; returns nnz */
int ngsx_csr_graph_finalize(ngsx_graph *g, int64_t *h_nnz, void *stream);
int ngsx_csr_graph_export(ngsx_graph *g, int64_t *rowptr /*[ndof+1]*/, int32_t *colidx /*[nnz]*/, void *stream);
int ngsx_csr_graph_destroy(ngsx_graph *g);
/* slots[k][i*nd+j] = position of (table[k][i], table[k][j]) in the CSR arrays (binary search of the sorted row) */
int ngsx_slot_map(int64_t n, int32_t nd, const int32_t *table, const int64_t *rowptr, const int32_t *colidx,
                  int32_t *slots, void *stream);
/* HOST function (all pointers host): greedy colouring of groups of `group` consecutive list entries such that two groups
 * of one colour share no dof (group == 1: NGSolve's element colouring).  h_table [n][nd] = dofs of list entry k.
 * Writes the colour-sorted permutation h_perm[n] (positions into the input list, groups kept contiguous) and
 * h_color_ptr[ncolors+1] (offsets into h_perm, in list entries).  cap_colors = capacity of h_color_ptr minus 1. */
int ngsx_color_groups(int64_t n, int32_t nd, const int32_t *h_table, int64_t ndof, int32_t group,
                      int64_t *h_perm, int64_t *h_color_ptr, int32_t cap_colors, int32_t *h_ncolors);
/* fused cut rule -> element matrix -> "+=" into CSR values, colour by colour, no floating-point atomics.
 * els/slots are in colour-sorted order (ngsx_color_groups), h_color_ptr is a HOST array. */
int ngsx_assemble_csr_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder,
                         const ngsx_form_h1 *h_form, int64_t nlist, const int32_t *els, const int32_t *slots,
                         int32_t ncolors, const int64_t *h_color_ptr, int32_t group, double *vals, void *stream);
int ngsx_assemble_vec_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder,
                         const ngsx_poly *h_f, int64_t nlist, const int32_t *els, const int32_t *el2dof_list /*[nlist][nd]*/,
                         int32_t ncolors, const int64_t *h_color_ptr, int32_t group, double *vec, void *stream);

/* ---- (d') owner-computes assembly: every CSR row is gathered by one warp, each value is written exactly once ------
 * Same result as ngsx_assemble_csr_h1 (SetZero + AddElementMatrix over all listed elements, NGS/comp/bilinearform.cpp:986,
 * 1563 + NGS/linalg/sparsematrix_impl.hpp:63-103) without colouring, zero-fill or read-modify-write: the plan holds, per
 * CSR row, the (element, local row) pairs contributing to it and the position of each local column inside the row; the
 * contributions of one entry are summed in list order by a single warp, so the values are bitwise reproducible and no
 * floating-point atomic is used.  Multi-GPU: a plan restricted to rows [row_lo, row_hi) makes the rank the owner of those
 * rows -- it recomputes the halo elements listed in `table` instead of exchanging interface rows.
 * table [nlist][nd]: dofs of list entry k (list order fixes the summation order), nd <= 26, nlist < 2^27, the nd dofs of one
 * entry distinct; CSR rows longer than 256 entries and dofs with more than 255 adjacent list entries are NGSX_ERR_UNSUPPORTED
 * (use the coloured path).  The plan is a static schedule (tiles of rows -> lanes -> steps, ngsxfem_b200/csrc/ngsx_rows.cuh). */
typedef struct ngsx_rowplan ngsx_rowplan;   /* opaque, owns device buffers (pair descriptors, geometry factors, cut-element matrices) */
int ngsx_rowplan_create(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                        int64_t row_lo, int64_t row_hi, ngsx_rowplan **out, void *stream);
/* Ownership by an arbitrary row set instead of a range: the plan covers the rows r with row_owner[r] == me (row_owner: device
 * int32 [ndof]).  Used by the multi-GPU partition: contiguous blocks of the ELEMENT list per rank, a row belongs to the rank of
 * the first listed element that touches it (ngsxfem_b200/partition.py) -- contiguous dof ranges would hand every rank all
 * vertex rows of an H1 space numbered "vertices, then edges".  The window of a tile may span rows of other owners; their
 * entries in the caller's (rank-local) value array are scratch. */
/* row plan over a lower-triangle pattern made with ngsx_csr_graph_set_lower: the local columns of a pair that lie above its row have
 * no slot and are skipped by the gather; every stored value is still written once, in plan order (bitwise the lower triangle of the
 * unsymmetric assembly).  Served by ngsx_assemble_csr_rows_h1 (rows of < 255 entries); the block gathers reject such plans. */
int ngsx_rowplan_create_lower(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                              int64_t row_lo, int64_t row_hi, ngsx_rowplan **out, void *stream);
int ngsx_rowplan_create_owned(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                              const int32_t *row_owner, int32_t me, ngsx_rowplan **out, void *stream);
/* h_info[0..5] = {non-empty rows, (row, element) pairs, CSR entries covered by the plan's rows, longest row, first covered CSR
 * entry, one past the last covered CSR entry} */
int ngsx_rowplan_info(const ngsx_rowplan *plan, int64_t *h_info);
int ngsx_rowplan_destroy(ngsx_rowplan *plan);
/* Cut-element buffer of ngsx_assemble_csr_rows_h1: sized by a counting pass on the plan's first numeric call (elements that
 * need quadrature + 25 %), reused afterwards.  ngsx_rowplan_cut_usage (synchronises the stream) returns the number of
 * quadrature elements the LAST call met and the capacity; count > capacity means that call skipped the excess elements
 * (the level set moved a lot): reserve more with ngsx_rowplan_reserve_cut and assemble again. */
int ngsx_rowplan_cut_usage(ngsx_rowplan *plan, int64_t *h_count, int64_t *h_capacity, void *stream);
int ngsx_rowplan_reserve_cut(ngsx_rowplan *plan, int64_t capacity);
/* els [nlist]: element numbers of the list entries (same order as `table` at plan creation).  Elements whose cut rule is
 * empty for (dt) contribute nothing (the reference's `ir == nullptr` return, xfem/symboliccutbfi.cpp:119-120).
 * accumulate == 0: rows of the plan are overwritten; != 0: added to.  err as in ngsx_classify (may be NULL). */
int ngsx_assemble_csr_rows_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder,
                              const ngsx_form_h1 *h_form, int64_t nlist, const int32_t *els, ngsx_rowplan *plan,
                              int accumulate, double *vals, int64_t *err, void *stream);

/* One-shot ordering hint for the NEXT ngsx_assemble_csr_rows_h1 call of this host thread: its stream waits for `event` (a
 * cudaEvent_t the caller recorded on another stream; NULL clears) after the cut-element stage and before the row gather.  The
 * gather is the one kernel that fills the machine; a caller with independent latency-bound work on side streams (CutInfo.Update,
 * Integrate: NGS task-parallel regions in the reference) overlaps it with the geometry / quadrature stage and keeps the
 * gather alone on the SMs.  Capturable (the wait becomes a graph edge). */
void ngsx_set_gather_wait_event(void *event);
/* General form of AddElementMatrix (NGS/linalg/sparsematrix_impl.hpp:63-103) over a plan: mats [nlist][nd][nd] row-major
 * blocks of the plan's list entries (elements, or facet patches with their merged dof lists) are gathered row by row;
 * accumulate as above.  nd in {3, 4, 5, 6, 8, 9, 10, 12, 14, 20}. */
int ngsx_assemble_csr_rows_mats(ngsx_rowplan *plan, const double *mats, int accumulate, double *vals, void *stream);
/* PACKED block rows: the fast form of the call above.  ngsx_rowplan_rowslots returns rowslot [nlist][nd] (device, owned by the
 * plan; built on the first request, synchronises): slot of the pair (list entry k, local row i) in the plan's schedule, -1 if the
 * row is not covered, and the number of slots.  A producer kernel writes row i of block k to rows + slot * ndp, ndp = nd rounded
 * up to even (ngsx_set_packed_output makes the block-producing entry points of this library do that); rows [nslots][ndp] must
 * be 16-byte aligned.  The rows a warp needs in one step are then contiguous: every byte is read once, in whole lines. */
int ngsx_rowplan_rowslots(ngsx_rowplan *plan, const int32_t **d_rowslot, int64_t *h_nslots, void *stream);
int ngsx_assemble_csr_rows_packed(ngsx_rowplan *plan, const double *rows, int accumulate, double *vals, void *stream);
/* One-shot, like ngsx_set_gather_wait_event: the NEXT block-producing call of this host thread (ngsx_elmat_h1_deformed,
 * ngsx_facetpatch_cmat_h1, ngsx_xfem_elmat_h1) writes its blocks as packed rows -- its output pointer is then the rows buffer
 * of ngsx_assemble_csr_rows_packed.  NULL clears. */
void ngsx_set_packed_output(const int32_t *d_rowslot);
/* The same for vectors (AddElementVector, NGS/comp/linearform.cpp:209-243, without atomics): out[d] (+)= scale * sum of
 * src[pairs[p]] over p in [dofptr[d], dofptr[d+1]), in list order.  src: flattened element vectors (e.g. of ngsx_elvec_h1),
 * pairs: indices into it grouped by dof -- entries a sign mask removes (DiffOpX, xfem/xfemdiffops.cpp:30-40) are simply not
 * listed.  Bitwise reproducible. */
int ngsx_gather_vec(int64_t ndof, const int64_t *dofptr /*[ndof+1]*/, const int64_t *pairs, const double *src, double scale,
                    int accumulate, double *out, void *stream);

/* ---- Integrate ------------------------------------------------------------------------------------------------- */
/* out[0] = sum_e sum_q |det J_e| w_q f(x_q) over the (dt, order) cut rule; fixed-shape tree reduction (bitwise
 * reproducible run to run).  out is a device double. */
int ngsx_integrate(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, const ngsx_poly *h_f, double *out, void *stream);

/* Allocation-free variant of ngsx_integrate.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, same
 * kernels and reduction tree, bitwise the result of ngsx_integrate; the scratch (per-warp partial sums, cut-element words) is a
 * file-scope __device__ array of the library sized for NGSX_STATIC_MAX_ELEMENTS elements (about 1.6 MB, held by every GPU
 * that loads the library until exit), so the call performs no cudaMalloc / cudaFree / memory-pool allocation.  Meshes with
 * ne > NGSX_STATIC_MAX_ELEMENTS are rejected with NGSX_ERR_CAPACITY.  All calls on one device share the array (it keeps the
 * last call's contents; nothing needs clearing): calls that overlap on the GPU -- different streams -- race, serialise them. */
#define NGSX_STATIC_MAX_ELEMENTS 4194304
int ngsx_integrate_static(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, const ngsx_poly *h_f, double *out, void *stream);

/* Fused sweep: ngsx_classify and up to 4 ngsx_integrate calls on the same level set in one pass over the element table
 * (CutInfo.Update followed by Integrate(...) calls is the usual sequence of a step; each of them alone streams el2v once and has
 * its own latency floor).  Outputs as ngsx_classify (any may be NULL); integral j over (h_dt[j], h_order[j]) of h_f[j] goes to
 * out[j] (device).  Bitwise the results of the separate calls: same partial sums, same reduction tree. */
int ngsx_classify_integrate(const ngsx_mesh *mesh, const double *lset_v, uint32_t *cdom_words, int8_t *dom, double *ratio, int64_t *err,
                            int32_t nint, const int32_t *h_dt, const int32_t *h_order, const ngsx_poly *h_f, double *out, void *stream);

/* ---- (f-4) deformed (isoparametric) elements -----------------------------------------------------------------------
 * mesh.SetDeformation(gf) (NGS/comp/bilinearform.cpp:1235 picks the deformed ElementTransformation; lsetcurving/lsetcurv.py
 * :108-163 produces gf): x(xi) = affine map + sum_i coef[el2dof[e][i]] phi_i(xi), phi = scalar H1 basis of `order` (1..2,
 * vertex dofs then edge dofs, as el2dof of the H1 space), coef interleaved [ndof][dim] (H1(mesh, order, dim=mesh.dim)).
 * The cut rule stays on the reference element (cutint/straightcutrule.cpp:580-661); measure, mapped gradients, interface
 * weight factor |F^-T n_ref| (:558-576), the level-set normal and h = specialcf.mesh_size = |det F|^(1/dim)
 * (NGS/fem/python_fem.cpp:564-593; not flagged elementwise constant, so xfem/symboliccutbfi.cpp:205-214 evaluates it at every
 * point) are evaluated per rule point. */
typedef struct ngsx_deformation {
  int32_t order;            /* 1 or 2 */
  int32_t nd;               /* dofs per element and component: dim+1 (order 1), 6 / 10 (order 2) */
  int64_t ndof;
  const int32_t *el2dof;    /* [ne][nd] */
  const double  *coef;      /* [ndof][dim] */
} ngsx_deformation;
/* batched CalcElementMatrixAdd (xfem/symboliccutbfi.cpp:73-315) on the deformed mesh; arguments as ngsx_elmat_h1, err as in
 * ngsx_classify (may be NULL).  The dense blocks reach the CSR values through ngsx_assemble_csr_rows_mats. */
int ngsx_elmat_h1_deformed(const ngsx_mesh *mesh, const ngsx_deformation *defo, const double *lset_v, int fe_order, int dt, int intorder,
                           const ngsx_form_h1 *h_form, int64_t nlist, const int32_t *els, double *elmats, int32_t *npts,
                           int64_t *err, void *stream);
/* batched CalcElementVector (xfem/symboliccutlfi.cpp:52-158) on the deformed mesh, f evaluated at the mapped points; arguments as
 * ngsx_elvec_h1.  The dense vectors reach the global vector through ngsx_gather_vec (fixed order, no atomics). */
int ngsx_elvec_h1_deformed(const ngsx_mesh *mesh, const ngsx_deformation *defo, const double *lset_v, int fe_order, int dt, int intorder,
                           const ngsx_poly *h_f, int64_t nlist, const int32_t *els, double *elvecs, int32_t *npts,
                           int64_t *err, void *stream);
/* IntegrateX (cutint/python_cutint.cpp:20-72) on the deformed mesh: f is evaluated at the mapped points */
int ngsx_integrate_deformed(const ngsx_mesh *mesh, const ngsx_deformation *defo, const double *lset_v, int dt, int order,
                            const ngsx_poly *h_f, double *out, void *stream);

/* ---- (f-3) space-time cut quadrature -----------------------------------------------------------------------------
 * SpaceTimeCutIntegrationRule (cutint/spacetimecutrule.cpp:98-196) for a level set that is P1 in space and nodal of order
 * ntime_dofs - 1 (0..5: NodalTimeFE, Gauss-Lobatto nodes, spacetime/SpaceTimeFE.cpp:150-173) in the reference time: lset_st [ntime_dofs][nv], time-major like SpaceTimeFESpace::GetDofNrs
 * (spacetime/SpaceTimeFESpace.cpp:109-122).  npts[e] = size of the element's space-time rule (0 where the reference returns
 * nullptr); points (reference coordinates), reference times and weights (the reference's wei_arr: w_space * w_time * (t1-t0),
 * IF weights carry |F^-T n|) are written at offs[e] in the reference's emission order.  Roots in time: closed forms for orders 1, 2
 * (spacetimecutrule.cpp:6-45), the reference's 50 subdivisions + secant / bisection refinement above (:46-95). */
int ngsx_st_rule_count(const ngsx_mesh *mesh, const double *lset_st, int ntime_dofs, int dt, int order_space, int order_time,
                       int32_t *npts, void *stream);
int ngsx_st_rule_fill(const ngsx_mesh *mesh, const double *lset_st, int ntime_dofs, int dt, int order_space, int order_time,
                      const int64_t *offs, double *pts3, double *tref, double *wts, void *stream);
/* IntegrateX with time_order >= 0 (cutint/python_cutint.cpp:20-72): out[0] = sum_e sum_q |det J_e| w_q f(x_q, tref_q),
 * f = sum_m coef[m] x^ex[m][0] y^ex[m][1] z^ex[m][2] tref^h_texp[m] (h_texp: host array [nterms] or NULL = no time dependence);
 * fixed-shape reduction, bitwise reproducible. */
int ngsx_st_integrate(const ngsx_mesh *mesh, const double *lset_st, int ntime_dofs, int dt, int order_space, int order_time,
                      const ngsx_poly *h_f, const int32_t *h_texp, double *out, void *stream);

/* Space-time element matrices: T_CalcElementMatrixAdd (xfem/symboliccutbfi.cpp:73-315) on the tensor-product element SpaceTimeFE
 * (spacetime/SpaceTimeFE.cpp:29-113: CalcShape, CalcDShape, CalcDtShape) = H1 of order fe_order (1..2) x NodalTimeFE with
 * fe_time_dofs dofs (1..6), dof j * nd_space + i, integrated with the space-time cut rule above:
 * elmats[k][(test)][(trial)] = sum_q |det J| w_q (c_mass u v + c_stiff grad u . grad v + c_dt dt(u) v), dt = derivative with respect
 * to the REFERENCE time (ngsxfem's dt(), python/__init__.py).  els: list of elements (NULL = all); blocks of elements with an empty
 * rule are zero. */
int ngsx_st_elmat_h1(const ngsx_mesh *mesh, const double *lset_st, int ntime_dofs, int fe_order, int fe_time_dofs, int dt, int order_space,
                     int order_time, double c_mass, double c_stiff, double c_dt, int64_t nlist, const int32_t *els, double *elmats, void *stream);
/* CutInformation::Update with time_order >= 0 (xfem/cutinfo.cpp:79-111): part volumes of the space-time prism from the POS and NEG
 * space-time rules of order 0 in space and time_order in time -> dom [ne] int8 (NEG / POS / IF), cut ratio [ne] */
int ngsx_st_classify(const ngsx_mesh *mesh, const double *lset_st, int ntime_dofs, int time_order, int8_t *dom, double *ratio, void *stream);

/* ---- (e) the exchange between the GPUs of one node -------------------------------------------------------------------
 * The reference's Integrate() sums over MPI ranks with a collective (cutint/python_cutint.cpp:66-70: AllReduce of the local sums).
 * Here each rank owns a mailbox in device memory, exported / mapped with CUDA IPC; ngsx_peer_exchange -- the last node of a rank's
 * captured step -- publishes the rank's n <= NGSX_PEER_MAXVALS local sums into every peer's mailbox with plain stores over NVLink
 * and collects the totals of the PREVIOUS step (out_prev, rank order: bitwise equal on all ranks); ngsx_peer_collect delivers the
 * totals of the last published step.  d_step: device counter, 0 before the first exchange, incremented by the kernel (replays of a
 * CUDA graph therefore advance it).  All ranks must call it once per step.  A peer that never publishes yields NaN, not a hang. */
#define NGSX_PEER_MAXVALS 4
int ngsx_peer_mailbox_bytes(int world);
int ngsx_peer_alloc(int world, void **mailbox, unsigned char *handle64 /* cudaIpcMemHandle_t, to be sent to the peers */);
int ngsx_peer_open(const unsigned char *handle64, void **mailbox);
int ngsx_peer_close(void *mailbox);
int ngsx_peer_free(void *mailbox);
int ngsx_peer_exchange(const double *local, int n, int rank, int world, void *const *d_peers, void *mailbox, int64_t *d_step, double *out_prev,
                       void *stream);
int ngsx_peer_collect(void *mailbox, int n, int world, const int64_t *d_step, double *out, void *stream);

/* ---- (f-5) quadrilaterals / hexahedra ------------------------------------------------------------------------------
 * LevelsetCutQuadrilateral (cutint/straightcutrule.cpp:206-501, policy FIND_OPTIMAL) and the ET_QUAD / ET_HEX branch of
 * StraightCutIntegrationRule[Untransformed] (:580-726).  For these entry points mesh->el2v is [ne][4] (dim 2) / [ne][8] (dim 3)
 * in NGSolve's local vertex order ((0,0), (1,0), (1,1), (0,1) [, the same at the upper face]); the level set is the multilinear
 * interpolant of the vertex values lset_v, the geometry the multilinear map of the vertices.  err as in ngsx_classify: elements
 * on which the reference throws (a slice that is not cut, an exact zero) -- and hexahedra that would take the reference's
 * fallback to tetrahedra (unreachable with a finite gradient) -- are counted, get no points and dom = -1.
 * ngsx_q_classify: CutInformation::Update (xfem/cutinfo.cpp:54-120): dom [ne] int8, cut ratio [ne] (order-0 part volumes).
 * ngsx_q_rule_count / _fill: as ngsx_rule_count / ngsx_rule_fill (reference coordinates, the reference's emission order; uncut
 * elements of the domain carry the tensor Gauss rule).  ngsx_q_integrate: IntegrateX (cutint/python_cutint.cpp:20-72). */
int ngsx_q_classify(const ngsx_mesh *mesh, const double *lset_v, int8_t *dom, double *ratio, int64_t *err, void *stream);
int ngsx_q_rule_count(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, int32_t *npts, int64_t *err, void *stream);
int ngsx_q_rule_fill(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, int transform_if, const int64_t *offs,
                     double *pts3, double *wts, void *stream);
int ngsx_q_integrate(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, const ngsx_poly *h_f, double *out, int64_t *err,
                     void *stream);

/* ---- facet / element / dof marking ----------------------------------------------------------------------------- */
/* facet2el [nf][2], -1 = no neighbour (boundary facet: bnd_val_a/b are used, xfem/cutinfo.cpp:238-241) */
int ngsx_facets_neighbor_types(int64_t nf, const int32_t *facet2el, const uint32_t *a_words, const uint32_t *b_words,
                               int bnd_val_a, int bnd_val_b, int use_and, uint32_t *out_words, void *stream);
/* the same on a mesh with periodic identifications (xfem/cutinfo.cpp:225-236): periodic_facet [nf] = MeshAccess::GetPeriodicFacet
 * (the partner facet, or the facet itself / -1 where there is none); a boundary facet with a partner of larger number takes the
 * partner's element as its second neighbour, the partner itself stays unmarked */
int ngsx_facets_neighbor_types_periodic(int64_t nf, const int32_t *facet2el, const int32_t *periodic_facet, const uint32_t *a_words,
                                        const uint32_t *b_words, int bnd_val_a, int bnd_val_b, int use_and, uint32_t *out_words, void *stream);
int ngsx_elements_with_neighbor_facets(int64_t nf, const int32_t *facet2el, const uint32_t *facet_words,
                                       int64_t ne, uint32_t *out_words /* zeroed by the callee */, void *stream);
int ngsx_dofs_of_elements(int64_t n, int32_t nd, const int32_t *table, const uint32_t *a_words,
                          int64_t ndof, uint32_t *out_words /* zeroed by the callee */, void *stream);

/* ---- ghost penalty: facet-patch matrices ------------------------------------------------------------------------ */
/* patch matrix of  coef * h^hpow * (u - u.Other()) (v - v.Other())  over both elements of each listed facet patch
 * (py_tutorials/fictdom_ghostpen.py:105-107); patch_els [np][2]; elmats [np][2nd][2nd] over [dofs(el0) | dofs(el1)]. */
int ngsx_facetpatch_elmat_h1(const ngsx_mesh *mesh, int fe_order, double coef, int hpow, int64_t npatch,
                             const int32_t *patch_els, double *elmats, void *stream);

/* ---- (f-2) XFEM: compound space [Vh, XFESpace(Vh)] with sign-masked operators --------------------------------------
 * One product term of a bilinear form on the compound space:
 *     coef * h^hpow * kappa * (L_v v) (L_u u)
 * kappa: 0 none, 1 CutRatioGF (cut ratio of the element, xfem/cutinfo.cpp:91), 2 one minus it;
 * pair : 0 value*value, 1 (test value)*(trial dn), 2 (test dn)*(trial value), 3 dn*dn, 4 grad.grad,  dn = grad . n,
 *        n = grad(lset_p1)/|grad(lset_p1)|;
 * blk_u / blk_v: which dofs carry the trial / test factor: 0 the dofs of Vh (u_std, grad(u_std)), 1 x-dofs of sign NEG
 *        (neg(u_x), neg_grad(u_x)), 2 x-dofs of sign POS (pos, pos_grad)  -- DiffOpX<D, RNEG/RPOS/RNEG_GRAD/RPOS_GRAD>,
 *        xfem/xfemdiffops.cpp:30-60. */
typedef struct ngsx_xterm {
  double  coef;
  int32_t hpow, kappa, pair, blk_u, blk_v, _pad;
} ngsx_xterm;
/* mats [nlist][2 nd][2 nd] row-major over [dofs of Vh | x-dofs] of each listed element (rows = test functions): the block
 * CalcElementMatrixAdd (xfem/symboliccutbfi.cpp:73-315) produces on a CompoundFiniteElement(H1 element, XFiniteElement).
 * xsign [nlist][nd] int8: XFESpace::domofdof (NEG/POS) of the x-dof over local dof i (xfem/xFESpace.cpp:346-366);
 * ratio [ne]: cut ratios of ngsx_classify.  Elements whose (dt) rule is empty give a zero block.  accumulate != 0: the block
 * is added to mats (several integrators of one form share one buffer and one gather). */
int ngsx_xfem_elmat_h1(const ngsx_mesh *mesh, const double *lset_v, int fe_order, int dt, int intorder, int32_t nterms,
                       const ngsx_xterm *h_terms, int64_t nlist, const int32_t *els, const int8_t *xsign, const double *ratio,
                       int accumulate, double *mats, void *stream);

/* the same on a deformed mesh (row f-4; DiffOpX evaluates CalcMappedDShape on the deformed ElementTransformation,
 * xfem/xfemdiffops.cpp:8-82): geometry, normal, interface weight factor and mesh_size per rule point */
int ngsx_xfem_elmat_h1_deformed(const ngsx_mesh *mesh, const ngsx_deformation *defo, const double *lset_v, int fe_order, int dt,
                                int intorder, int32_t nterms, const ngsx_xterm *h_terms, int64_t nlist, const int32_t *els,
                                const int8_t *xsign, const double *ratio, int accumulate, double *mats, void *stream);

/* Merged patch blocks for the row gather (NGSolve's compressed_dnums / compressed_elmat, NGS/comp/bilinearform.cpp:1800-1853).
 * ngsx_facetpatch_table_h1 (symbolic, synchronises): table [np][ndp] = dofs of element 0, then the dofs of element 1 that
 * element 0 does not have; cidx [np][2 nd] = position of every local dof of [dofs(el0) | dofs(el1)] in that list.
 * ndp = 2 nd - (dofs on the shared facet): 4 / 9 (trig P1 / P2), 5 / 14 (tet P1 / P2).
 * ngsx_facetpatch_cmat_h1 (numeric): mats [np][ndp][ndp] = C em C^T of ngsx_facetpatch_elmat_h1, integrated in merged form. */
int ngsx_facetpatch_table_h1(int64_t npatch, const int32_t *patch_els, int32_t nd, const int32_t *el2dof, int32_t ndp,
                             int32_t *table, uint8_t *cidx, void *stream);
int ngsx_facetpatch_cmat_h1(const ngsx_mesh *mesh, int fe_order, double coef, int hpow, int64_t npatch,
                            const int32_t *patch_els, const uint8_t *cidx, int32_t ndp, double *mats, void *stream);

/* the same on a deformed mesh (row f-4): a rule point of one element is pulled into the other through both isoparametric maps by
 * the Newton iteration of MapPatchIntegrationPoint (xfem/symboliccutbfi.cpp:1113-1199: affine start, <= 200 steps, tolerance
 * 1e-12 sqrt(det F), "low order guess" exit); weights mir1.GetWeight(), h = mesh_size on mir1 per point (:1359-1365) */
int ngsx_facetpatch_cmat_h1_deformed(const ngsx_mesh *mesh, const ngsx_deformation *defo, int fe_order, double coef, int hpow,
                                     int64_t npatch, const int32_t *patch_els, const uint8_t *cidx, int32_t ndp, double *mats,
                                     void *stream);

#ifdef __cplusplus
}
#endif
#endif /* NGSX_ABI_H */
