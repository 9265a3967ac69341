// ngsx_rows.cu -- (d') owner-computes assembly: plan construction and the row-gather kernel.
//
// The coloured scatter (ngsx_assemble.cu) streams the whole CSR value array once per colour: on the 10M-tet config every
// 32-byte sector of the 584 MB value array is read and written by ~13 different launches.  Here the dependency is turned
// around: the plan lists, for every non-empty CSR row, the (element, local row) pairs that contribute to it, and ONE WARP
// gathers a row completely -- contributions are summed in list order in shared memory and each value leaves the SM once,
// as a coalesced store.  Constant-rule (uncut) elements are recomputed on the fly from 7 geometry factors per element
// (56 B, L2-resident) contracted with per-CTA reference tables; cut elements read their row of a compact element-matrix
// buffer produced by the quadrature kernel.  No colouring, no zero-fill, no read-modify-write, no floating-point atomics.
//
// Work decomposition of the row kernel: warp = tile of consecutive rows (values in the warp's shared-memory window), lane =
// contiguous range of the tile's rows (see k_rows).
// HBM traffic per P2 tet pair: 16 B descriptor + (L2) 64 B factors; per CSR entry: 8 B written once.
#include <cub/cub.cuh>
#include <vector>
#include <thrust/iterator/counting_iterator.h>
#include "ngsx_rows.cuh"

namespace ngsx {

// ---- plan construction --------------------------------------------------------------------------------------------
// key = (row, local row i, list index k): the pairs of a row are walked grouped by i, so that the lanes of a warp mostly
// read the same rows of the reference tables in the same step (shared-memory broadcast instead of distinct wavefronts)
__global__ void __launch_bounds__(256) k_pair_keys(int64_t npairs, int nd, int64_t nlist, const int32_t *__restrict__ table, int64_t row_lo,
                                                  int64_t row_hi, const int32_t *__restrict__ row_owner, int32_t me, uint64_t *__restrict__ keys) {
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npairs) return;
  int32_t r = table[p];
  const int64_t k = p / nd, i = p - k * nd;
  bool mine = r >= row_lo && r < row_hi;
  if (mine && row_owner) mine = row_owner[r] == me;
  keys[p] = mine ? (((uint64_t)(uint32_t)r << 32) | (uint64_t)(i * nlist + k)) : ~0ULL;
}

__global__ void k_lower_bound(int64_t n, const uint64_t *__restrict__ keys, uint64_t target, int64_t *out) {
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (keys[mid] < target) lo = mid + 1; else hi = mid;
  }
  *out = lo;
}

__global__ void __launch_bounds__(256) k_head_flags(int64_t n, const uint64_t *__restrict__ keys, uint8_t *__restrict__ flags) {
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  flags[p] = (p == 0) || ((keys[p] >> 32) != (keys[p - 1] >> 32));
}

// rows[q] = row of the first pair of run q; closes pairptr; row statistics
__global__ void __launch_bounds__(256) k_plan_rows(int64_t nrows, int64_t npairs, const uint64_t *__restrict__ keys, int64_t *__restrict__ pairptr,
                                                  const int64_t *__restrict__ rowptr, int32_t *__restrict__ rows,
                                                  unsigned long long *__restrict__ stat /* {covered, maxlen, lo, hi} */) {
  int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q == 0) pairptr[nrows] = npairs;
  if (q >= nrows) return;
  int32_t r = (int32_t)(keys[pairptr[q]] >> 32);
  rows[q] = r;
  unsigned long long len = (unsigned long long)(rowptr[r + 1] - rowptr[r]);
  atomicAdd(&stat[0], len);
  atomicMax(&stat[1], len);
  if (q == 0) stat[2] = (unsigned long long)rowptr[r];
  if (q == nrows - 1) stat[3] = (unsigned long long)rowptr[r + 1];
}

// one thread per (pair, local column): position of the column inside the CSR row
__global__ void __launch_bounds__(256) k_plan_desc(int64_t npairs, int nd, int64_t nlist, int ds, const uint64_t *__restrict__ keys,
                                                  const int32_t *__restrict__ table, const int64_t *__restrict__ rowptr,
                                                  const int32_t *__restrict__ colidx, uint8_t *__restrict__ desc, int *__restrict__ bad) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= npairs * nd) return;
  int64_t pp = t / nd;
  int j = (int)(t - pp * nd);
  uint64_t key = keys[pp];
  int32_t r = (int32_t)(key >> 32);
  uint32_t p = (uint32_t)(key & 0xffffffffULL);
  int i = (int)(p / (uint32_t)nlist);
  uint32_t k = p - (uint32_t)i * (uint32_t)nlist;
  int32_t col = table[(int64_t)k * nd + j];
  int64_t rs = rowptr[r], lo = rs, hi = rowptr[r + 1] - 1;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (colidx[mid] < col) lo = mid + 1; else hi = mid;
  }
  int64_t pos = lo - rs;
  if (hi < rs || colidx[lo] != col) { atomicOr(bad, 1); pos = 0; }
  if (pos >= MAXROW) { atomicOr(bad, 2); pos = 0; }
  uint8_t *d = desc + pp * ds;
  if (j == 0) { *reinterpret_cast<int32_t *>(d) = (int32_t)k; d[4] = (uint8_t)i; }
  d[5 + j] = (uint8_t)pos;
}

__global__ void __launch_bounds__(256) k_row_spans(int64_t nrows, const int32_t *__restrict__ rows, const int64_t *__restrict__ rowptr,
                                                  int64_t *__restrict__ span) {
  int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nrows) return;
  const int32_t r = rows[q];
  span[2 * q] = rowptr[r]; span[2 * q + 1] = rowptr[r + 1];
}

// processing order of the tiles: by the list index of the first contributing element, so that the vertex-row and edge-row
// tiles of one mesh region run at about the same time (their geometry factors are then L2 hits); empty tiles sort last
__global__ void __launch_bounds__(256) k_tile_keys(int64_t ntiles, const int32_t *__restrict__ tile_q, const int64_t *__restrict__ pairptr,
                                                  const uint8_t *__restrict__ desc, int ds, uint32_t *__restrict__ keys, int32_t *__restrict__ ids,
                                                  unsigned long long *__restrict__ nonempty) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ntiles) return;
  int q0 = tile_q[c], q1 = tile_q[c + 1];
  ids[c] = (int32_t)c;
  if (q0 == q1) { keys[c] = 0xffffffffu; return; }
  keys[c] = (uint32_t)*reinterpret_cast<const int32_t *>(desc + pairptr[q0] * ds);
  atomicAdd(nonempty, 1ULL);
}

// per tile, in processing order: CSR span and the row range of each lane -- the tile's pairs are cut into 32 equal parts
// and every cut is moved to the next row boundary
__global__ void __launch_bounds__(256) k_tile_tables(int64_t nact, const int32_t *__restrict__ order, const int32_t *__restrict__ tile_q,
                                                    const int32_t *__restrict__ rows, const int64_t *__restrict__ rowptr,
                                                    const int64_t *__restrict__ pairptr, TileInfo *__restrict__ info, int32_t *__restrict__ split) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nact * 33) return;
  const int64_t tt = t / 33;
  const int k = (int)(t - tt * 33);
  const int c = order[tt];
  const int q0 = tile_q[c], q1 = tile_q[c + 1];
  if (k == 32) {
    TileInfo ti;
    ti.z0 = rowptr[rows[q0]];
    ti.zlen = (int32_t)(rowptr[rows[q1 - 1] + 1] - ti.z0);
    ti.pad = 0;
    info[tt] = ti;
    split[t] = q1;
    return;
  }
  const int64_t pa = pairptr[q0], pb = pairptr[q1];
  const int64_t target = pa + ((pb - pa) * k) / 32;
  int lo = q0, hi = q1;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (pairptr[mid] < target) lo = mid + 1; else hi = mid;
  }
  split[t] = lo;
}

// byte 5 + j of a descriptor = position of local column j in the row
template <int J> __device__ __forceinline__ int desc_pos(const uint4 &d) {
  constexpr int b = 5 + J, w = b >> 2, sh = (b & 3) * 8;
  const uint32_t x = w == 1 ? d.y : (w == 2 ? d.z : d.w);
  return (int)((x >> sh) & 0xffu);
}

// positions of the nd local columns (compile-time byte extraction)
template <int ND, int J> __device__ __forceinline__ void desc_positions(const uint4 &d, int *pos) {
  if constexpr (J < ND) { pos[J] = desc_pos<J>(d); desc_positions<ND, J + 1>(d, pos); }
}

// 32-byte global load (sm_100: LDG.E.256): one L1 tag lookup per half geometry record instead of two
__device__ __forceinline__ void ldg256(const double *p, double *o) {
  asm volatile("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(o[0]), "=d"(o[1]), "=d"(o[2]), "=d"(o[3]) : "l"(p));
}

// Row i of a constant-rule element matrix from the element's geometry factors, by sum factorisation.
//   A_ij = |J| c_s sum_r w_r (grad phi_i)^T (J^-1 J^-T) (grad phi_j) + |J| c_m sum_r w_r phi_i phi_j
// The reference gradients of P1/P2 shapes are affine in the reference coordinates, d_a phi_i = sum_m c_iam X_m, so with the
// rule's moments folded into cM[i][a][n] (k_ref_tables):  w[b][n] = sum_a Ghat_ab cM[i][a][n]  (36 FMAs in 3D) and
//   vertex j:      A_ij = w[.][0] . grad lam_j
//   edge (a, b):   A_ij = -1/2 (Q[a][b] + Q[b][a]),  Q[a][k] = grad lam_a . W(., lam_k),  W(b, lam_k) = coefficient of lam_k in w[b][.]
// Same quadrature as the reference's static rule, re-associated; 12 + nd doubles of table per pair instead of 7 nd.
template <int DIM, int ORDER>
__device__ __forceinline__ void row_values(const double *g, const double *cmi, const double *tm, double *v) {
  constexpr int ND = H1<DIM, ORDER>::ND, NEDGE = H1<DIM, ORDER>::NEDGE;
  double Gh[DIM][DIM];
  {
    int pr = 0;
#pragma unroll
    for (int a = 0; a < DIM; a++)
#pragma unroll
      for (int b = a; b < DIM; b++) { Gh[a][b] = g[pr]; Gh[b][a] = g[pr]; pr++; }
  }
  double c[DIM][DIM + 1];
  if constexpr (DIM == 3) {
#pragma unroll
    for (int t = 0; t < 6; t++) {
      const double2 x = *reinterpret_cast<const double2 *>(cmi + 2 * t);
      c[(2 * t) / 4][(2 * t) % 4] = x.x; c[(2 * t + 1) / 4][(2 * t + 1) % 4] = x.y;
    }
  } else {
#pragma unroll
    for (int t = 0; t < 3; t++) {
      const double2 x = *reinterpret_cast<const double2 *>(cmi + 2 * t);
      c[(2 * t) / 3][(2 * t) % 3] = x.x; c[(2 * t + 1) / 3][(2 * t + 1) % 3] = x.y;
    }
  }
  double w[DIM][DIM + 1];
#pragma unroll
  for (int b = 0; b < DIM; b++)
#pragma unroll
    for (int n = 0; n <= DIM; n++) {
      double s = Gh[0][b] * c[0][n];
#pragma unroll
      for (int a = 1; a < DIM; a++) s = fma(Gh[a][b], c[a][n], s);
      w[b][n] = s;
    }
  // vertex functions: grad lam_j = e_j (j < DIM), -(1,..,1) (j = DIM)
  double dsum = 0.0;
#pragma unroll
  for (int j = 0; j < DIM; j++) { v[j] = w[j][0]; dsum += w[j][0]; }
  v[DIM] = -dsum;
  if constexpr (ORDER == 2) {
    double Q[DIM + 1][DIM + 1];
#pragma unroll
    for (int b = 0; b < DIM; b++) {
      double last = w[b][0];
#pragma unroll
      for (int k = 0; k < DIM; k++) { Q[b][k] = w[b][k + 1]; last -= w[b][k + 1]; }
      Q[b][DIM] = last;
    }
#pragma unroll
    for (int k = 0; k <= DIM; k++) {
      double z = Q[0][k];
#pragma unroll
      for (int b = 1; b < DIM; b++) z += Q[b][k];
      Q[DIM][k] = -z;
    }
#pragma unroll
    for (int ed = 0; ed < NEDGE; ed++) {
      int a, b;
      edge_vertices<DIM>(ed, a, b);
      v[DIM + 1 + ed] = -0.5 * (Q[a][b] + Q[b][a]);
    }
  }
  // mass
  const double gm = g[6];
  if constexpr (ND % 2 == 0) {
#pragma unroll
    for (int j = 0; j < ND; j += 2) {
      const double2 t = *reinterpret_cast<const double2 *>(tm + j);
      v[j] = fma(gm, t.x, v[j]); v[j + 1] = fma(gm, t.y, v[j + 1]);
    }
  } else {
#pragma unroll
    for (int j = 0; j < ND; j++) v[j] = fma(gm, tm[j], v[j]);
  }
}

// row i of a quadrature (cut) element matrix: nd contiguous doubles of the cut buffer
template <int ND>
__device__ __forceinline__ void row_values_cut(const double *cm, double *v) {
  if constexpr (ND % 2 == 0) {
#pragma unroll
    for (int j = 0; j < ND; j += 2) {
      const double2 t = __ldg(reinterpret_cast<const double2 *>(cm + j));
      v[j] = t.x; v[j + 1] = t.y;
    }
  } else {
#pragma unroll
    for (int j = 0; j < ND; j++) v[j] = __ldg(cm + j);
  }
}

// ---- the row kernel ---------------------------------------------------------------------------------------------
// WARP = one tile of consecutive CSR rows whose values (<= ACCW doubles) live in the warp's private shared-memory window:
// initialised by the warp (zero, or the current values when accumulating), gathered, and streamed out once with coalesced
// stores.  Warps never wait for each other (no CTA barrier after the table load).
// LANE = one contiguous range of the tile's rows (pair-balanced split from the plan): the lane walks the (row, element)
// pairs of its range in plan order; per pair it contracts the element's 7 geometry factors with row i of the reference
// tables (nd entries, shared-memory vector loads) and adds them at their positions of the row.  No two lanes share a row
// and the nd columns of one element are distinct, so the summation order of every CSR entry is the plan order: bitwise
// reproducible, no atomics.  Descriptors are prefetched two pairs ahead, geometry factors one pair ahead, row records one
// row ahead; the start-up loads of the next tile are issued before the current tile is streamed out.
struct RowArgs {
  int64_t ntiles;             // non-empty tiles, in processing order
  const TileInfo *info;       // [ntiles]
  const int32_t *split;       // [ntiles][33] first plan row of each lane's range
  const RowRec *rowrec;
  const uint4 *desc;
  const double *G;
  const double *cutmats;
  const double *tab;
  double *vals;
  int accumulate;
};

// per-lane walk state over the pairs [pcur, plast) of the rows [q, qb)
struct LaneState {
  int q, qb, off;
  uint32_t pcur, pend, plast;
  RowRec recn;
  uint4 d1, d2;
  double g1[8];               // geometry record of pair pcur (g1[7] = code)
};

__device__ __forceinline__ void lane_start(const RowArgs &A, int64_t tt, int lane, LaneState &S) {
  const int32_t *split = A.split + tt * 33;
  S.q = __ldg(split + lane); S.qb = __ldg(split + lane + 1);
  S.pcur = S.pend = S.plast = 0; S.off = 0;
  S.recn = RowRec{0, 0, 0};
  S.d1 = S.d2 = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
  for (int m = 0; m < 8; m++) S.g1[m] = 0.0;
  if (S.q < S.qb) {
    const RowRec rec = ld_rowrec(A.rowrec + S.q), recl = ld_rowrec(A.rowrec + S.qb - 1);
    S.pcur = rec.p0; S.pend = rec.p0 + rec.np; S.off = rec.off; S.plast = recl.p0 + recl.np;
    S.recn = (S.q + 1 < S.qb) ? ld_rowrec(A.rowrec + S.q + 1) : rec;
    S.d1 = __ldg(A.desc + S.pcur);
    if (S.pcur + 1 < S.plast) S.d2 = __ldg(A.desc + S.pcur + 1);
    const double *Gk = A.G + (int64_t)(int)S.d1.x * GSTRIDE;
    ldg256(Gk, S.g1); ldg256(Gk + 4, S.g1 + 4);
  }
}

template <int DIM, int ORDER>
__global__ void __launch_bounds__(RW * 32, NGSX_MINBLOCKS) k_rows(RowArgs A) {
  constexpr int ND = H1<DIM, ORDER>::ND, NE = ND * ND, CS = RefTab<DIM>::CS, CMO = RefTab<DIM>::cm_offset(NE), NTAB = CMO + ND * CS;
  extern __shared__ __align__(16) double smem_rows[];
  double *tab = smem_rows;                                        // mass table [ND][ND], then cM [ND][CS]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *acc = smem_rows + NTAB + warp * ACCW;                    // this warp's window
  for (int t = threadIdx.x; t < NTAB; t += blockDim.x) tab[t] = A.tab[t];
  __syncthreads();
  const int64_t nwarps = (int64_t)gridDim.x * RW;
  int64_t tt = (int64_t)blockIdx.x * RW + warp;
  LaneState S;
  TileInfo ti{0, 0, 0};
  if (tt < A.ntiles) { ti = A.info[tt]; lane_start(A, tt, lane, S); }
  while (tt < A.ntiles) {
    const int64_t z0 = ti.z0;
    const int zlen = ti.zlen;
    if (A.accumulate) for (int t = lane; t < zlen; t += 32) acc[t] = A.vals[z0 + t];
    else for (int t = lane; t < zlen; t += 32) acc[t] = 0.0;
    __syncwarp();
    while (S.pcur < S.plast) {
      const uint4 d = S.d1;
      double g[8];
#pragma unroll
      for (int m = 0; m < 8; m++) g[m] = S.g1[m];
      const long long code = __double_as_longlong(g[7]);
      if (S.pcur == S.pend) {                      // first pair of the next row
        S.q++;
        S.pend = S.recn.p0 + S.recn.np; S.off = S.recn.off;
        if (S.q + 1 < S.qb) S.recn = ld_rowrec(A.rowrec + S.q + 1);
      }
      if (S.pcur + 1 < S.plast) {                  // advance the prefetch pipeline
        S.d1 = S.d2;
        const double *Gk = A.G + (int64_t)(int)S.d1.x * GSTRIDE;
        ldg256(Gk, S.g1); ldg256(Gk + 4, S.g1 + 4);
        if (S.pcur + 2 < S.plast) S.d2 = __ldg(A.desc + S.pcur + 2);
      }
      if (code >= 0) {
        const int i = (int)(d.y & 0xffu);
        double v[ND], a[ND];
        int pos[ND];
        desc_positions<ND, 0>(d, pos);
        double *accrow = acc + S.off;
        // the nd columns of an element are distinct: read all, add, write all (no serialised read-modify-write chain)
#pragma unroll
        for (int j = 0; j < ND; j++) a[j] = accrow[pos[j]];
        if (code > 0) row_values_cut<ND>(A.cutmats + (code - 1) * NE + i * ND, v);
        else row_values<DIM, ORDER>(g, tab + CMO + i * CS, tab + i * ND, v);
#pragma unroll
        for (int j = 0; j < ND; j++) accrow[pos[j]] = a[j] + v[j];
      }
      S.pcur++;
    }
    // next tile: issue its start-up loads now, they fly while this tile is streamed out
    tt += nwarps;
    if (tt < A.ntiles) { ti = A.info[tt]; lane_start(A, tt, lane, S); }
    __syncwarp();
    for (int t = lane; t < zlen; t += 32) A.vals[z0 + t] = acc[t];
    __syncwarp();
  }
}

template <int DIM, int ORDER>
static int launch_rows_t(const RowArgs &A, int sm_count, cudaStream_t st) {
  constexpr int ND = H1<DIM, ORDER>::ND, NTAB = RefTab<DIM>::cm_offset(ND * ND) + ND * RefTab<DIM>::CS;
  size_t smem = (size_t)(NTAB + RW * ACCW) * sizeof(double);
  auto kern = k_rows<DIM, ORDER>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, RW * 32, smem));
  if (occ < 1) { set_error("row kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
  int64_t need = (A.ntiles + RW - 1) / RW;
  unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)sm_count * occ));
  kern<<<grid, RW * 32, smem, st>>>(A);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int rows_scratch(ngsx_rowplan *P, int npairtab) {
  const int64_t ne2 = (int64_t)P->nd * P->nd;
  if (!P->G) NGSX_CUDA(cudaMalloc(&P->G, std::max<int64_t>(1, P->nlist) * GSTRIDE * sizeof(double)));
  if (!P->cutlist) NGSX_CUDA(cudaMalloc(&P->cutlist, std::max<int64_t>(1, P->nlist) * sizeof(int32_t)));
  if (!P->cutcount) NGSX_CUDA(cudaMalloc(&P->cutcount, sizeof(int32_t)));
  if (!P->cutmats) NGSX_CUDA(cudaMalloc(&P->cutmats, std::max<int64_t>(1, P->nlist) * ne2 * sizeof(double)));
  if (!P->tab) NGSX_CUDA(cudaMalloc(&P->tab, (size_t)(npairtab + 1) * ne2 * sizeof(double)));
  return NGSX_OK;
}

int launch_rows(const ngsx_rowplan *P, int npairtab, int accumulate, double *vals, int sm_count, cudaStream_t st) {
  if (P->nrows == 0) return NGSX_OK;
  if (P->ds != 16) { set_error("row gather: this plan was built for merged facet-patch dof lists (use ngsx_assemble_csr_rows_mats)"); return NGSX_ERR_ARG; }
  RowArgs A{P->ntiles_active, P->tile_info, P->tile_split, P->rowrec, reinterpret_cast<const uint4 *>(P->desc), P->G, P->cutmats, P->tab,
            vals, accumulate};
  switch (P->nd * 10 + npairtab) {
    case 33: return launch_rows_t<2, 1>(A, sm_count, st);
    case 63: return launch_rows_t<2, 2>(A, sm_count, st);
    case 46: return launch_rows_t<3, 1>(A, sm_count, st);
    case 106: return launch_rows_t<3, 2>(A, sm_count, st);
  }
  set_error("row gather: unsupported (dofs per element, dimension)");
  return NGSX_ERR_UNSUPPORTED;
}

}  // namespace ngsx

using namespace ngsx;

extern "C" {

int ngsx_rowplan_destroy(ngsx_rowplan *P) {
  if (!P) return NGSX_OK;
  cudaDeviceSynchronize();
  cudaFree(P->rows); cudaFree(P->pairptr); cudaFree(P->desc); cudaFree(P->rowrec); cudaFree(P->tile_q); cudaFree(P->tile_info); cudaFree(P->tile_split);
  cudaFree(P->G); cudaFree(P->cutlist); cudaFree(P->cutcount); cudaFree(P->cutmats); cudaFree(P->tab);
  delete P;
  return NGSX_OK;
}

int ngsx_rowplan_info(const ngsx_rowplan *P, int64_t *h_info) {
  NGSX_REQUIRE(P && h_info, "rowplan_info: bad arguments");
  h_info[0] = P->nrows; h_info[1] = P->npairs; h_info[2] = P->nnz_covered; h_info[3] = P->max_row;
  h_info[4] = P->nnz_lo; h_info[5] = P->nnz_hi;
  return NGSX_OK;
}

static int rowplan_create(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                          int64_t row_lo, int64_t row_hi, const int32_t *row_owner, int32_t me, ngsx_rowplan **out, void *stream);

int ngsx_rowplan_create(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                        int64_t row_lo, int64_t row_hi, ngsx_rowplan **out, void *stream) {
  return rowplan_create(nlist, nd, table, ndof, rowptr, colidx, row_lo, row_hi, nullptr, 0, out, stream);
}

int ngsx_rowplan_create_owned(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                              const int32_t *row_owner, int32_t me, ngsx_rowplan **out, void *stream) {
  if (!row_owner) { set_error("rowplan_create_owned: row_owner is NULL"); return NGSX_ERR_ARG; }
  return rowplan_create(nlist, nd, table, ndof, rowptr, colidx, 0, -1, row_owner, me, out, stream);
}

static int rowplan_create(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                          int64_t row_lo, int64_t row_hi, const int32_t *row_owner, int32_t me, ngsx_rowplan **out, void *stream) {
  NGSX_REQUIRE(out && nlist >= 0 && nd > 0 && nd <= 27 && ndof >= 0 && ndof < (1LL << 31) && rowptr, "rowplan_create: bad arguments");
  NGSX_REQUIRE(nlist == 0 || (table && colidx), "rowplan_create: table / colidx are NULL");
  NGSX_REQUIRE(nlist * nd < (1LL << 32), "rowplan_create: more than 2^32 (element, dof) pairs");
  cudaStream_t st = as_stream(stream);
  if (row_hi < 0 || row_hi > ndof) row_hi = ndof;
  if (row_lo < 0) row_lo = 0;
  ngsx_rowplan *P = new ngsx_rowplan();
  P->nlist = nlist; P->nd = nd; P->ndof = ndof; P->ds = nd <= 11 ? 16 : 32; P->rowptr = rowptr;
  *out = P;
  const int64_t nall = nlist * nd;
  if (nall == 0 || row_hi <= row_lo) return NGSX_OK;

  uint64_t *keys = nullptr, *alt = nullptr;
  uint8_t *flags = nullptr;
  void *tmp = nullptr;
  int64_t *d_scalar = nullptr;          // [0] valid pairs, [1] rows
  unsigned long long *d_stat = nullptr;
  int *d_bad = nullptr;
  auto cleanup = [&]() { cudaFree(keys); cudaFree(alt); cudaFree(flags); cudaFree(tmp); cudaFree(d_scalar); cudaFree(d_stat); cudaFree(d_bad); };
#define PLAN_CUDA(call)                                                                                         \
  do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { set_error(std::string(#call) + ": " + cudaGetErrorString(e__)); \
       cleanup(); ngsx_rowplan_destroy(P); *out = nullptr; return NGSX_ERR_CUDA; } } while (0)

  PLAN_CUDA(cudaMalloc(&keys, nall * sizeof(uint64_t)));
  PLAN_CUDA(cudaMalloc(&alt, nall * sizeof(uint64_t)));
  PLAN_CUDA(cudaMalloc(&d_scalar, 2 * sizeof(int64_t)));
  PLAN_CUDA(cudaMalloc(&d_stat, 4 * sizeof(unsigned long long)));
  PLAN_CUDA(cudaMalloc(&d_bad, sizeof(int)));
  PLAN_CUDA(cudaMemsetAsync(d_stat, 0, 4 * sizeof(unsigned long long), st));
  PLAN_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
  k_pair_keys<<<(unsigned)((nall + 255) / 256), 256, 0, st>>>(nall, nd, nlist, table, row_lo, row_hi, row_owner, me, keys);
  cub::DoubleBuffer<uint64_t> db(keys, alt);
  size_t tmp_bytes = 0;
  PLAN_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, db, nall, 0, 64, st));
  PLAN_CUDA(cudaMalloc(&tmp, tmp_bytes));
  PLAN_CUDA(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, db, nall, 0, 64, st));
  const uint64_t *sorted = db.Current();
  k_lower_bound<<<1, 1, 0, st>>>(nall, sorted, (uint64_t)row_hi << 32, d_scalar);
  int64_t npairs = 0;
  PLAN_CUDA(cudaMemcpyAsync(&npairs, d_scalar, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaStreamSynchronize(st));
  g_launches.fetch_add(4);
  P->npairs = npairs;
  if (npairs == 0) { cleanup(); return NGSX_OK; }

  PLAN_CUDA(cudaMalloc(&flags, npairs));
  k_head_flags<<<(unsigned)((npairs + 255) / 256), 256, 0, st>>>(npairs, sorted, flags);
  // pair offsets of the row heads: at most min(npairs, rows in range) of them
  int64_t cap_rows = std::min<int64_t>(npairs, row_hi - row_lo);
  PLAN_CUDA(cudaMalloc(&P->pairptr, (cap_rows + 1) * sizeof(int64_t)));
  cudaFree(tmp); tmp = nullptr; tmp_bytes = 0;
  thrust::counting_iterator<int64_t> counting(0);
  PLAN_CUDA(cub::DeviceSelect::Flagged(nullptr, tmp_bytes, counting, flags, P->pairptr, d_scalar + 1, npairs, st));
  PLAN_CUDA(cudaMalloc(&tmp, tmp_bytes));
  PLAN_CUDA(cub::DeviceSelect::Flagged(tmp, tmp_bytes, counting, flags, P->pairptr, d_scalar + 1, npairs, st));
  int64_t nrows = 0;
  PLAN_CUDA(cudaMemcpyAsync(&nrows, d_scalar + 1, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaStreamSynchronize(st));
  g_launches.fetch_add(3);
  P->nrows = nrows;
  PLAN_CUDA(cudaMalloc(&P->rows, nrows * sizeof(int32_t)));
  PLAN_CUDA(cudaMalloc(&P->desc, (size_t)npairs * P->ds));
  PLAN_CUDA(cudaMemsetAsync(P->desc, 0, (size_t)npairs * P->ds, st));
  k_plan_rows<<<(unsigned)((nrows + 255) / 256), 256, 0, st>>>(nrows, npairs, sorted, P->pairptr, rowptr, P->rows, d_stat);
  k_plan_desc<<<(unsigned)((npairs * nd + 255) / 256), 256, 0, st>>>(npairs, nd, nlist, P->ds, sorted, table, rowptr, colidx, P->desc, d_bad);
  g_launches.fetch_add(2);
  unsigned long long stat[4];
  int bad = 0;
  PLAN_CUDA(cudaMemcpyAsync(stat, d_stat, sizeof stat, cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof bad, cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaStreamSynchronize(st));
  PLAN_CUDA(cudaGetLastError());
  P->nnz_covered = (int64_t)stat[0]; P->max_row = (int64_t)stat[1]; P->nnz_lo = (int64_t)stat[2]; P->nnz_hi = (int64_t)stat[3];
  if (!(bad & 3) && P->max_row <= MAXROW) {
    // Tiles: greedy over the plan rows, on the host (a sequential scan of 8 bytes per row; symbolic phase).  A tile is closed
    // when its CSR span would exceed the warp's window, at TILE_MAXROWS rows, or at TILE_MAXPAIRS pairs -- so tiles of long
    // rows (24 pairs on a P2 vertex row) hold 32 rows, one per lane, and tiles of short rows two per lane.
    std::vector<int64_t> h_span(2 * nrows), h_pairptr(nrows + 1);
    {
      int64_t *d_span = nullptr;
      PLAN_CUDA(cudaMalloc(&d_span, 2 * nrows * sizeof(int64_t)));
      k_row_spans<<<(unsigned)((nrows + 255) / 256), 256, 0, st>>>(nrows, P->rows, rowptr, d_span);
      cudaError_t ce = cudaMemcpyAsync(h_span.data(), d_span, 2 * nrows * sizeof(int64_t), cudaMemcpyDeviceToHost, st);
      if (ce == cudaSuccess) ce = cudaMemcpyAsync(h_pairptr.data(), P->pairptr, (nrows + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, st);
      if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);
      cudaFree(d_span);
      PLAN_CUDA(ce);
    }
    std::vector<int32_t> h_tile_q;
    std::vector<RowRec> h_rowrec(nrows);
    h_tile_q.reserve(nrows / 16 + 2);
    int64_t q0 = 0;
    h_tile_q.push_back(0);
    for (int64_t q = 0; q < nrows; q++) {
      const int64_t np = h_pairptr[q + 1] - h_pairptr[q];
      if (np > 65535) bad |= 4;
      if (q > q0 && (h_span[2 * q + 1] - h_span[2 * q0] > ACCW || q - q0 >= TILE_MAXROWS ||
                     h_pairptr[q + 1] - h_pairptr[q0] > TILE_MAXPAIRS)) {
        q0 = q;
        h_tile_q.push_back((int32_t)q);
      }
      RowRec r;
      r.p0 = (uint32_t)h_pairptr[q]; r.np = (uint16_t)np; r.off = (uint16_t)(h_span[2 * q] - h_span[2 * q0]);
      h_rowrec[q] = r;
    }
    h_tile_q.push_back((int32_t)nrows);
    P->ntiles = (int64_t)h_tile_q.size() - 1;
    PLAN_CUDA(cudaMalloc(&P->tile_q, (P->ntiles + 1) * sizeof(int32_t)));
    PLAN_CUDA(cudaMalloc(&P->rowrec, nrows * sizeof(RowRec)));
    PLAN_CUDA(cudaMemcpyAsync(P->tile_q, h_tile_q.data(), (P->ntiles + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    PLAN_CUDA(cudaMemcpyAsync(P->rowrec, h_rowrec.data(), nrows * sizeof(RowRec), cudaMemcpyHostToDevice, st));
    PLAN_CUDA(cudaStreamSynchronize(st));          // the host vectors go out of use only after the copies
    g_launches.fetch_add(1);
    // tile processing order
    uint32_t *ck = nullptr, *ck2 = nullptr;
    int32_t *ci2 = nullptr, *order = nullptr;
    PLAN_CUDA(cudaMalloc(&ck, P->ntiles * sizeof(uint32_t)));
    PLAN_CUDA(cudaMalloc(&ck2, P->ntiles * sizeof(uint32_t)));
    PLAN_CUDA(cudaMalloc(&ci2, P->ntiles * sizeof(int32_t)));
    PLAN_CUDA(cudaMalloc(&order, P->ntiles * sizeof(int32_t)));
    PLAN_CUDA(cudaMemsetAsync(d_stat, 0, sizeof(unsigned long long), st));
    k_tile_keys<<<(unsigned)((P->ntiles + 255) / 256), 256, 0, st>>>(P->ntiles, P->tile_q, P->pairptr, P->desc, P->ds, ck, ci2, d_stat);
    cudaFree(tmp); tmp = nullptr; tmp_bytes = 0;
    PLAN_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, ck, ck2, ci2, order, (int)P->ntiles, 0, 32, st));
    PLAN_CUDA(cudaMalloc(&tmp, tmp_bytes));
    PLAN_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, ck, ck2, ci2, order, (int)P->ntiles, 0, 32, st));
    g_launches.fetch_add(5);
    unsigned long long nonempty = 0;
    PLAN_CUDA(cudaMemcpyAsync(&nonempty, d_stat, sizeof nonempty, cudaMemcpyDeviceToHost, st));
    PLAN_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof bad, cudaMemcpyDeviceToHost, st));
    PLAN_CUDA(cudaStreamSynchronize(st));
    P->ntiles_active = (int64_t)nonempty;
    PLAN_CUDA(cudaMalloc(&P->tile_info, std::max<int64_t>(1, P->ntiles_active) * sizeof(TileInfo)));
    PLAN_CUDA(cudaMalloc(&P->tile_split, std::max<int64_t>(1, P->ntiles_active) * 33 * sizeof(int32_t)));
    if (P->ntiles_active > 0)
      k_tile_tables<<<(unsigned)((P->ntiles_active * 33 + 255) / 256), 256, 0, st>>>(P->ntiles_active, order, P->tile_q, P->rows, rowptr, P->pairptr,
                                                                                   P->tile_info, P->tile_split);
    g_launches.fetch_add(1);
    PLAN_CUDA(cudaStreamSynchronize(st));
    cudaFree(ck); cudaFree(ck2); cudaFree(ci2); cudaFree(order);
    PLAN_CUDA(cudaGetLastError());
  }
#undef PLAN_CUDA
  cleanup();
  if (bad & 1) {
    set_error("rowplan_create: an element coupling is missing from the sparsity pattern (element_restriction too small?)");
    ngsx_rowplan_destroy(P); *out = nullptr;
    return NGSX_ERR_ARG;
  }
  if (bad & 4) {
    set_error("rowplan_create: a dof with more than 65535 adjacent elements");
    ngsx_rowplan_destroy(P); *out = nullptr;
    return NGSX_ERR_UNSUPPORTED;
  }
  if ((bad & 2) || P->max_row > MAXROW) {
    set_error("rowplan_create: a CSR row is longer than 256 entries; use the coloured scatter (ngsx_assemble_csr_h1)");
    ngsx_rowplan_destroy(P); *out = nullptr;
    return NGSX_ERR_UNSUPPORTED;
  }
  return NGSX_OK;
}

}  // extern "C"
