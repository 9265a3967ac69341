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
// Work decomposition of the row kernel: CTA = chunk of consecutive rows (values in shared memory), lane group of nd lanes =
// one row at a time, lane j = column j of the pair's element matrix (see k_rows).
// HBM traffic per P2 tet pair: 16 B descriptor + (L2) 64 B factors; per CSR entry: 8 B written once.
#include <cub/cub.cuh>
#include <thrust/iterator/counting_iterator.h>
#include "ngsx_rows.cuh"

namespace ngsx {

constexpr int RWARPS = 8;
constexpr int MAXROW = 256;       // longest supported CSR row (column positions are bytes)
constexpr int CHW = 2816;         // chunk window in CSR entries: a chunk holds the rows that START inside one window
constexpr int ACC = CHW + MAXROW; // shared-memory accumulator of a chunk (24 KB)

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

// chunk c = plan rows whose first CSR entry lies in [nnz_lo + c CHW, nnz_lo + (c+1) CHW): first such row by binary search
__global__ void __launch_bounds__(256) k_plan_chunks(int64_t nchunks, int64_t nrows, int64_t nnz_lo, const int32_t *__restrict__ rows,
                                                    const int64_t *__restrict__ rowptr, int32_t *__restrict__ chunk_q) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c > nchunks) return;
  if (c == nchunks) { chunk_q[c] = (int32_t)nrows; return; }
  const int64_t target = nnz_lo + c * CHW;
  int64_t lo = 0, hi = nrows;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (rowptr[rows[mid]] < target) lo = mid + 1; else hi = mid;
  }
  chunk_q[c] = (int32_t)lo;
}

__global__ void __launch_bounds__(256) k_plan_rowrec(int64_t nrows, int64_t nnz_lo, const int32_t *__restrict__ rows, const int64_t *__restrict__ pairptr,
                                                    const int64_t *__restrict__ rowptr, const int32_t *__restrict__ chunk_q,
                                                    RowRec *__restrict__ rowrec, int *__restrict__ bad) {
  int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nrows) return;
  const int64_t zr = rowptr[rows[q]];
  const int64_t c = (zr - nnz_lo) / CHW;
  const int64_t z0 = rowptr[rows[chunk_q[c]]];
  const int64_t np = pairptr[q + 1] - pairptr[q];
  if (np > 65535) atomicOr(bad, 4);
  RowRec r;
  r.p0 = (uint32_t)pairptr[q]; r.np = (uint16_t)np; r.off = (uint16_t)(zr - z0);
  rowrec[q] = r;
}

// ---- the row kernel ---------------------------------------------------------------------------------------------
// CTA = one chunk of consecutive CSR rows whose values (<= ACC doubles) live in shared memory: initialised cooperatively
// (zero, or the current values when accumulating), gathered, and streamed out once with coalesced stores.
// Inside the chunk every lane group ("slot", nd lanes) owns whole rows: slot s walks rows q0+s, q0+s+NSLOT, ... pair by pair,
// lane j of the slot computing entry (i, j) of the pair's element matrix and adding it at its position of the row.
// Slots never share a row, so there is no ordering problem between them; consecutive pairs of one row are ordered by the
// per-step __syncwarp().  Summation order of every CSR entry = list order of the plan: bitwise reproducible.
struct RowArgs {
  int64_t nchunks;
  const int32_t *chunk_q;     // [nchunks + 1] first plan row of each chunk
  const int32_t *rows;
  const RowRec *rowrec;
  const uint4 *desc;
  const int64_t *rowptr;
  const double *G;
  const double *cutmats;
  const double *tab;
  double *vals;
  int accumulate;
};

template <int ND, int NP>
__global__ void __launch_bounds__(RWARPS * 32) k_rows(RowArgs A) {
  constexpr int NE = ND * ND, PG = 32 / ND, NSLOT = RWARPS * PG;
  extern __shared__ double smem_rows[];
  double *tab = smem_rows;                       // [(NP+1)][NE]
  double *acc = smem_rows + (NP + 1) * NE;        // [ACC] values of the chunk's rows
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int t = threadIdx.x; t < (NP + 1) * NE; t += blockDim.x) tab[t] = A.tab[t];
  const int g = lane / ND, j = lane - g * ND;
  const bool lane_on = g < PG;
  const int slot = warp * PG + g;
  const int wsel = (5 + j) >> 2, wsh = ((5 + j) & 3) * 8;   // descriptor word / shift of pos[j]
  for (int64_t c = blockIdx.x; c < A.nchunks; c += gridDim.x) {
    const int q0 = __ldg(A.chunk_q + c), q1 = __ldg(A.chunk_q + c + 1);
    if (q0 == q1) continue;
    const int64_t z0 = __ldg(A.rowptr + __ldg(A.rows + q0));
    const int zlen = (int)(__ldg(A.rowptr + __ldg(A.rows + q1 - 1) + 1) - z0);
    __syncthreads();                               // tab loaded / previous chunk streamed out
    if (A.accumulate) for (int t = threadIdx.x; t < zlen; t += blockDim.x) acc[t] = A.vals[z0 + t];
    else for (int t = threadIdx.x; t < zlen; t += blockDim.x) acc[t] = 0.0;
    __syncthreads();
    int q = q0 + slot;
    uint32_t pcur = 0, pend = 0;
    int off = 0;
    for (;;) {
      if (lane_on && pcur == pend && q < q1) {
        const RowRec rec = A.rowrec[q];
        pcur = rec.p0; pend = rec.p0 + rec.np; off = rec.off;
        q += NSLOT;
      }
      const bool work = lane_on && pcur < pend;
      if (!__any_sync(0xffffffffu, work)) break;
      if (work) {
        const uint4 d = __ldg(A.desc + pcur);
        const int k = (int)d.x, i = (int)(d.y & 0xffu);
        const uint32_t w = wsel == 1 ? d.y : (wsel == 2 ? d.z : d.w);
        const int pos = (int)((w >> wsh) & 0xffu);
        const double2 *Gk = reinterpret_cast<const double2 *>(A.G + (int64_t)k * GSTRIDE);
        const double2 g3 = __ldg(Gk + 3);
        const long long code = __double_as_longlong(g3.y);
        const int ij = i * ND + j;
        if (code == 0) {
          const double2 g0 = __ldg(Gk), g1 = __ldg(Gk + 1);
          const double *t = tab + ij;
          double v;
          if constexpr (NP == 6) {
            const double2 g2 = __ldg(Gk + 2);
            v = g0.x * t[0] + g0.y * t[NE] + g1.x * t[2 * NE] + g1.y * t[3 * NE] + g2.x * t[4 * NE] + g2.y * t[5 * NE] + g3.x * t[6 * NE];
          } else {
            v = g0.x * t[0] + g0.y * t[NE] + g1.x * t[2 * NE] + g3.x * t[3 * NE];
          }
          acc[off + pos] += v;
        } else if (code > 0) {
          acc[off + pos] += __ldg(A.cutmats + (code - 1) * NE + ij);
        }
        pcur++;
      }
      __syncwarp();
    }
    __syncthreads();
    for (int t = threadIdx.x; t < zlen; t += blockDim.x) A.vals[z0 + t] = acc[t];
  }
}

template <int ND, int NP>
static int launch_rows_t(const RowArgs &A, int sm_count, cudaStream_t st) {
  constexpr int NE = ND * ND;
  size_t smem = (size_t)((NP + 1) * NE + ACC) * sizeof(double);
  auto kern = k_rows<ND, NP>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, RWARPS * 32, smem));
  if (occ < 1) { set_error("row kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
  unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(A.nchunks, (int64_t)sm_count * occ));
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
  RowArgs A{P->nchunks, P->chunk_q, P->rows, P->rowrec, reinterpret_cast<const uint4 *>(P->desc), P->rowptr, P->G, P->cutmats, P->tab,
            vals, accumulate};
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
  cudaFree(P->rows); cudaFree(P->pairptr); cudaFree(P->desc); cudaFree(P->rowrec); cudaFree(P->chunk_q);
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
  NGSX_REQUIRE(out && nlist >= 0 && nd > 0 && nd <= 11 && ndof >= 0 && ndof < (1LL << 31) && rowptr, "rowplan_create: bad arguments");
  NGSX_REQUIRE(nlist == 0 || (table && colidx), "rowplan_create: table / colidx are NULL");
  NGSX_REQUIRE(nlist * nd < (1LL << 32), "rowplan_create: more than 2^32 (element, dof) pairs");
  cudaStream_t st = as_stream(stream);
  if (row_hi < 0 || row_hi > ndof) row_hi = ndof;
  if (row_lo < 0) row_lo = 0;
  ngsx_rowplan *P = new ngsx_rowplan();
  P->nlist = nlist; P->nd = nd; P->ndof = ndof; P->ds = 16; P->rowptr = rowptr;
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
  P->nnz_covered = (int64_t)stat[0]; P->max_row = (int64_t)stat[1]; P->nnz_lo = (int64_t)stat[2]; P->nnz_hi = (int64_t)stat[3];
  if (!(bad & 3) && P->max_row <= MAXROW) {
    P->nchunks = (P->nnz_hi - P->nnz_lo + CHW - 1) / CHW;
    PLAN_CUDA(cudaMalloc(&P->chunk_q, (P->nchunks + 1) * sizeof(int32_t)));
    PLAN_CUDA(cudaMalloc(&P->rowrec, nrows * sizeof(RowRec)));
    k_plan_chunks<<<(unsigned)((P->nchunks + 1 + 255) / 256), 256, 0, st>>>(P->nchunks, nrows, P->nnz_lo, P->rows, rowptr, P->chunk_q);
    k_plan_rowrec<<<(unsigned)((nrows + 255) / 256), 256, 0, st>>>(nrows, P->nnz_lo, P->rows, P->pairptr, rowptr, P->chunk_q, P->rowrec, d_bad);
    g_launches.fetch_add(2);
    PLAN_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof bad, cudaMemcpyDeviceToHost, st));
    PLAN_CUDA(cudaStreamSynchronize(st));
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
