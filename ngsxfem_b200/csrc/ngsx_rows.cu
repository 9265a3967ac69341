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
// Lane layout of the row kernel: a warp step handles PG = 32/nd pairs, lane = g*nd + j computes entry (i_g, j) of pair g;
// the PG contributions are added to the row accumulator in PG ordered sub-steps (two pairs of one row usually share columns).
// HBM traffic per P2 tet pair: 16 B descriptor + (L2) 64 B factors; per CSR entry: 8 B written once.
#include <cub/cub.cuh>
#include <thrust/iterator/counting_iterator.h>
#include "ngsx_rows.cuh"

namespace ngsx {

constexpr int RWARPS = 8;
constexpr int MAXROW = 256;

// ---- plan construction --------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_pair_keys(int64_t npairs, const int32_t *__restrict__ table, int64_t row_lo, int64_t row_hi,
                                                  uint64_t *__restrict__ keys) {
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npairs) return;
  int32_t r = table[p];
  keys[p] = (r < row_lo || r >= row_hi) ? ~0ULL : (((uint64_t)(uint32_t)r << 32) | (uint64_t)p);
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
__global__ void __launch_bounds__(256) k_plan_desc(int64_t npairs, int nd, int ds, const uint64_t *__restrict__ keys,
                                                  const int32_t *__restrict__ table, const int64_t *__restrict__ rowptr,
                                                  const int32_t *__restrict__ colidx, uint8_t *__restrict__ desc, int *__restrict__ bad) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= npairs * nd) return;
  int64_t pp = t / nd;
  int j = (int)(t - pp * nd);
  uint64_t key = keys[pp];
  int32_t r = (int32_t)(key >> 32);
  uint32_t p = (uint32_t)(key & 0xffffffffULL);
  uint32_t k = p / (uint32_t)nd;
  int i = (int)(p - k * (uint32_t)nd);
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

// ---- the row kernel ---------------------------------------------------------------------------------------------
struct RowArgs {
  int64_t nrows;
  const int32_t *rows;
  const int64_t *pairptr;
  const uint8_t *desc;
  const int64_t *rowptr;
  const double *G;
  const double *cutmats;
  const double *tab;
  double *vals;
  int accumulate;
};

template <int ND, int NP>
__global__ void __launch_bounds__(RWARPS * 32) k_rows(RowArgs A) {
  constexpr int NE = ND * ND, PG = 32 / ND, DS = (5 + ND + 3) & ~3;
  extern __shared__ double smem_rows[];
  double *tab = smem_rows;                                  // [(NP+1)][NE]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *acc = smem_rows + (NP + 1) * NE + warp * MAXROW;   // this warp's row accumulator
  for (int t = threadIdx.x; t < (NP + 1) * NE; t += blockDim.x) tab[t] = A.tab[t];
  __syncthreads();
  const int g = lane / ND, j = lane - g * ND;
  const bool lane_on = g < PG;
  for (int64_t rq = (int64_t)blockIdx.x * RWARPS + warp; rq < A.nrows; rq += (int64_t)gridDim.x * RWARPS) {
    const int32_t r = __ldg(A.rows + rq);
    const int64_t p0 = __ldg(A.pairptr + rq), p1 = __ldg(A.pairptr + rq + 1);
    const int64_t rs = __ldg(A.rowptr + r);
    const int len = (int)(__ldg(A.rowptr + r + 1) - rs);
    for (int c = lane; c < len; c += 32) acc[c] = 0.0;
    __syncwarp();
    for (int64_t p = p0; p < p1; p += PG) {
      const int64_t pp = p + g;
      double v = 0.0;
      int pos = 0;
      bool add = false;
      if (lane_on && pp < p1) {
        const uint8_t *d = A.desc + pp * DS;
        const int k = *reinterpret_cast<const int32_t *>(d);
        const int i = d[4];
        pos = d[5 + j];
        const double2 *Gk = reinterpret_cast<const double2 *>(A.G + (int64_t)k * GSTRIDE);
        const double2 g3 = __ldg(Gk + 3);
        const long long code = __double_as_longlong(g3.y);
        const int ij = i * ND + j;
        if (code == 0) {
          const double2 g0 = __ldg(Gk), g1 = __ldg(Gk + 1);
          const double *t = tab + ij;
          if constexpr (NP == 6) {
            const double2 g2 = __ldg(Gk + 2);
            v = g0.x * t[0] + g0.y * t[NE] + g1.x * t[2 * NE] + g1.y * t[3 * NE] + g2.x * t[4 * NE] + g2.y * t[5 * NE] + g3.x * t[6 * NE];
          } else {
            v = g0.x * t[0] + g0.y * t[NE] + g1.x * t[2 * NE] + g3.x * t[3 * NE];
          }
          add = true;
        } else if (code > 0) {
          v = __ldg(A.cutmats + (code - 1) * NE + ij);
          add = true;
        }
      }
#pragma unroll
      for (int s = 0; s < PG; s++) {
        if (add && g == s) acc[pos] += v;
        __syncwarp();
      }
    }
    if (A.accumulate) {
      for (int c = lane; c < len; c += 32) A.vals[rs + c] += acc[c];
    } else {
      for (int c = lane; c < len; c += 32) A.vals[rs + c] = acc[c];
    }
    __syncwarp();
  }
}

template <int ND, int NP>
static int launch_rows_t(const RowArgs &A, int sm_count, cudaStream_t st) {
  constexpr int NE = ND * ND;
  size_t smem = (size_t)((NP + 1) * NE + RWARPS * MAXROW) * sizeof(double);
  auto kern = k_rows<ND, NP>;
  int occ = 0;
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, RWARPS * 32, smem));
  if (occ < 1) { set_error("row kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
  int64_t need = (A.nrows + RWARPS - 1) / RWARPS;
  unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)sm_count * occ));
  kern<<<grid, RWARPS * 32, smem, st>>>(A);
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
  RowArgs A{P->nrows, P->rows, P->pairptr, P->desc, P->rowptr, P->G, P->cutmats, P->tab, vals, accumulate};
  switch (P->nd * 10 + npairtab) {
    case 33: return launch_rows_t<3, 3>(A, sm_count, st);
    case 63: return launch_rows_t<6, 3>(A, sm_count, st);
    case 46: return launch_rows_t<4, 6>(A, sm_count, st);
    case 106: return launch_rows_t<10, 6>(A, sm_count, st);
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
  cudaFree(P->rows); cudaFree(P->pairptr); cudaFree(P->desc);
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

int ngsx_rowplan_create(int64_t nlist, int32_t nd, const int32_t *table, int64_t ndof, const int64_t *rowptr, const int32_t *colidx,
                        int64_t row_lo, int64_t row_hi, ngsx_rowplan **out, void *stream) {
  NGSX_REQUIRE(out && nlist >= 0 && nd > 0 && nd <= 27 && ndof >= 0 && ndof < (1LL << 31) && rowptr, "rowplan_create: bad arguments");
  NGSX_REQUIRE(nlist == 0 || (table && colidx), "rowplan_create: table / colidx are NULL");
  NGSX_REQUIRE(nlist * nd < (1LL << 32), "rowplan_create: more than 2^32 (element, dof) pairs");
  cudaStream_t st = as_stream(stream);
  if (row_hi < 0 || row_hi > ndof) row_hi = ndof;
  if (row_lo < 0) row_lo = 0;
  ngsx_rowplan *P = new ngsx_rowplan();
  P->nlist = nlist; P->nd = nd; P->ndof = ndof; P->ds = (5 + nd + 3) & ~3; P->rowptr = rowptr;
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
  k_pair_keys<<<(unsigned)((nall + 255) / 256), 256, 0, st>>>(nall, table, row_lo, row_hi, keys);
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
  k_plan_desc<<<(unsigned)((npairs * nd + 255) / 256), 256, 0, st>>>(npairs, nd, P->ds, sorted, table, rowptr, colidx, P->desc, d_bad);
  g_launches.fetch_add(2);
  unsigned long long stat[4];
  int bad = 0;
  PLAN_CUDA(cudaMemcpyAsync(stat, d_stat, sizeof stat, cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof bad, cudaMemcpyDeviceToHost, st));
  PLAN_CUDA(cudaStreamSynchronize(st));
  PLAN_CUDA(cudaGetLastError());
#undef PLAN_CUDA
  cleanup();
  P->nnz_covered = (int64_t)stat[0]; P->max_row = (int64_t)stat[1]; P->nnz_lo = (int64_t)stat[2]; P->nnz_hi = (int64_t)stat[3];
  if (bad & 1) {
    set_error("rowplan_create: an element coupling is missing from the sparsity pattern (element_restriction too small?)");
    ngsx_rowplan_destroy(P); *out = nullptr;
    return NGSX_ERR_ARG;
  }
  if ((bad & 2) || P->max_row > MAXROW) {
    set_error("rowplan_create: a CSR row is longer than 256 entries; use the coloured scatter (ngsx_assemble_csr_h1)");
    ngsx_rowplan_destroy(P); *out = nullptr;
    return NGSX_ERR_UNSUPPORTED;
  }
  return NGSX_OK;
}

}  // extern "C"
