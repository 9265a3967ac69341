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
// Work decomposition of the row kernel: CTA = chunk of consecutive rows (values in shared memory), one thread per
// contiguous range of rows (see k_rows).
// HBM traffic per P2 tet pair: 16 B descriptor + (L2) 64 B factors; per CSR entry: 8 B written once.
#include <cub/cub.cuh>
#include <thrust/iterator/counting_iterator.h>
#include "ngsx_rows.cuh"

namespace ngsx {

constexpr int RT = 256;           // threads per CTA of the row kernel
constexpr int MAXROW = 256;       // longest supported CSR row (column positions are bytes)
constexpr int CHW = 7936;         // chunk window in CSR entries: a chunk holds the rows that START inside one window
constexpr int ACC = CHW + MAXROW; // shared-memory accumulator of a chunk (64 KB)

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

// thread t of the CTA that owns chunk c walks rows [split[t], split[t+1]): the chunk's pairs are cut into RT equal parts
// and every cut is moved to the next row boundary
__global__ void __launch_bounds__(256) k_chunk_split(int64_t nchunks, const int32_t *__restrict__ chunk_q, const int64_t *__restrict__ pairptr,
                                                    int32_t *__restrict__ split) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nchunks * (RT + 1)) return;
  const int64_t c = t / (RT + 1);
  const int k = (int)(t - c * (RT + 1));
  const int q0 = chunk_q[c], q1 = chunk_q[c + 1];
  if (k == RT || q0 == q1) { split[t] = k == RT ? q1 : q0; return; }
  const int64_t pa = pairptr[q0], pb = pairptr[q1];
  const int64_t target = pa + ((pb - pa) * k) / RT;
  int lo = q0, hi = q1;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (pairptr[mid] < target) lo = mid + 1; else hi = mid;
  }
  split[t] = lo;
}

// processing order of the chunks: by the list index of the first contributing element, so that the vertex-row and edge-row
// chunks of one mesh region run at about the same time; empty chunks get the largest key
__global__ void __launch_bounds__(256) k_chunk_keys(int64_t nchunks, const int32_t *__restrict__ chunk_q, const int64_t *__restrict__ pairptr,
                                                   const uint8_t *__restrict__ desc, int ds, uint32_t *__restrict__ keys, int32_t *__restrict__ ids,
                                                   unsigned long long *__restrict__ nonempty) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nchunks) return;
  int q0 = chunk_q[c], q1 = chunk_q[c + 1];
  ids[c] = (int32_t)c;
  if (q0 == q1) { keys[c] = 0xffffffffu; return; }
  keys[c] = (uint32_t)*reinterpret_cast<const int32_t *>(desc + pairptr[q0] * ds);
  atomicAdd(nonempty, 1ULL);
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
// ONE THREAD PER ROW RANGE: the rows of the chunk are split into RT contiguous ranges, a thread walks the (row, element)
// pairs of its range in plan order; per pair it contracts the element's 7 geometry factors with row i of the reference
// tables (nd entries, shared-memory vector loads) and adds them at their positions of the row.  No two threads share a
// row, so no synchronisation is needed between the cooperative init and the stream-out, and the summation order of every
// CSR entry is the plan order: bitwise reproducible.  Descriptors are prefetched two pairs ahead, geometry factors one
// pair ahead, row records one row ahead.
struct RowArgs {
  int64_t nchunks;            // non-empty chunks
  const int32_t *chunk_order; // [nchunks] chunk numbers in processing order (mesh locality: geometry factors stay in L2)
  const int32_t *chunk_q;     // [.. + 1] first plan row of each chunk
  const int32_t *chunk_split; // [..][RT + 1] first plan row of each thread's range
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

__device__ __forceinline__ RowRec ld_rowrec(const RowRec *p) {
  const uint2 u = __ldg(reinterpret_cast<const uint2 *>(p));
  RowRec r;
  r.p0 = u.x; r.np = (uint16_t)(u.y & 0xffffu); r.off = (uint16_t)(u.y >> 16);
  return r;
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

// row i of the pair's element matrix: nd values, either contracted from the geometry factors or read from the cut buffer
template <int ND, int NP>
__device__ __forceinline__ void row_values(const double *gq, const double *tb, const double *cm, bool iscut, double *v) {
  constexpr int NE = ND * ND;
  if (iscut) {
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
    return;
  }
#pragma unroll
  for (int j = 0; j < ND; j++) v[j] = 0.0;
#pragma unroll
  for (int q = 0; q <= NP; q++) {
    if constexpr (ND % 2 == 0) {
#pragma unroll
      for (int j = 0; j < ND; j += 2) {
        const double2 t = *reinterpret_cast<const double2 *>(tb + q * NE + j);
        v[j] += gq[q] * t.x; v[j + 1] += gq[q] * t.y;
      }
    } else {
#pragma unroll
      for (int j = 0; j < ND; j++) v[j] += gq[q] * tb[q * NE + j];
    }
  }
}

template <int ND, int NP>
__global__ void __launch_bounds__(RT) k_rows(RowArgs A) {
  constexpr int NE = ND * ND;
  extern __shared__ __align__(16) double smem_rows[];
  double *tab = smem_rows;                       // [(NP+1)][NE]
  double *acc = smem_rows + (NP + 1) * NE;        // [ACC] values of the chunk's rows
  for (int t = threadIdx.x; t < (NP + 1) * NE; t += blockDim.x) tab[t] = A.tab[t];
  for (int64_t cc = blockIdx.x; cc < A.nchunks; cc += gridDim.x) {
    const int c = __ldg(A.chunk_order + cc);
    const int q0 = __ldg(A.chunk_q + c), q1 = __ldg(A.chunk_q + c + 1);
    const int64_t z0 = __ldg(A.rowptr + __ldg(A.rows + q0));
    const int zlen = (int)(__ldg(A.rowptr + __ldg(A.rows + q1 - 1) + 1) - z0);
    // this thread's rows [qa, qb) (pair-balanced split from the plan) and pairs [pcur, plast)
    const int32_t *split = A.chunk_split + (int64_t)c * (RT + 1);
    const int qa = __ldg(split + threadIdx.x), qb = __ldg(split + threadIdx.x + 1);
    int q = qa;
    uint32_t pcur = 0, pend = 0, plast = 0;
    int off = 0;
    RowRec recn{0, 0, 0};
    uint4 d1 = make_uint4(0u, 0u, 0u, 0u), d2 = d1;
    double2 g1[4];
    g1[0] = g1[1] = g1[2] = g1[3] = make_double2(0.0, 0.0);
    if (qa < qb) {
      const RowRec rec = ld_rowrec(A.rowrec + qa), recl = ld_rowrec(A.rowrec + qb - 1);
      pcur = rec.p0; pend = rec.p0 + rec.np; off = rec.off; plast = recl.p0 + recl.np;
      recn = (qa + 1 < qb) ? ld_rowrec(A.rowrec + qa + 1) : rec;
      d1 = __ldg(A.desc + pcur);
      if (pcur + 1 < plast) d2 = __ldg(A.desc + pcur + 1);
      const double2 *Gk = reinterpret_cast<const double2 *>(A.G + (int64_t)(int)d1.x * GSTRIDE);
#pragma unroll
      for (int m = 0; m < 4; m++) g1[m] = __ldg(Gk + m);
    }
    __syncthreads();                               // tab loaded / previous chunk streamed out
    if (A.accumulate) for (int t = threadIdx.x; t < zlen; t += blockDim.x) acc[t] = A.vals[z0 + t];
    else for (int t = threadIdx.x; t < zlen; t += blockDim.x) acc[t] = 0.0;
    __syncthreads();
    while (pcur < plast) {
      const uint4 d = d1;
      double gq[NP + 1];
      if constexpr (NP == 6) { gq[0] = g1[0].x; gq[1] = g1[0].y; gq[2] = g1[1].x; gq[3] = g1[1].y; gq[4] = g1[2].x; gq[5] = g1[2].y; gq[6] = g1[3].x; }
      else { gq[0] = g1[0].x; gq[1] = g1[0].y; gq[2] = g1[1].x; gq[3] = g1[3].x; }
      const long long code = __double_as_longlong(g1[3].y);
      if (pcur == pend) {                          // first pair of the next row
        q++;
        pend = recn.p0 + recn.np; off = recn.off;
        if (q + 1 < qb) recn = ld_rowrec(A.rowrec + q + 1);
      }
      if (pcur + 1 < plast) {                      // advance the prefetch pipeline
        d1 = d2;
        const double2 *Gk = reinterpret_cast<const double2 *>(A.G + (int64_t)(int)d1.x * GSTRIDE);
#pragma unroll
        for (int m = 0; m < 4; m++) g1[m] = __ldg(Gk + m);
        if (pcur + 2 < plast) d2 = __ldg(A.desc + pcur + 2);
      }
      if (code >= 0) {
        const int i = (int)(d.y & 0xffu);
        const bool iscut = code > 0;
        const double *cm = A.cutmats + (iscut ? (code - 1) * NE + i * ND : 0);
        double v[ND], a[ND];
        int pos[ND];
        desc_positions<ND, 0>(d, pos);
        double *accrow = acc + off;
        // the nd columns of an element are distinct: read all, add, write all (no serialised read-modify-write chain)
#pragma unroll
        for (int j = 0; j < ND; j++) a[j] = accrow[pos[j]];
        row_values<ND, NP>(gq, tab + i * ND, cm, iscut, v);
#pragma unroll
        for (int j = 0; j < ND; j++) accrow[pos[j]] = a[j] + v[j];
      }
      pcur++;
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
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, RT, smem));
  if (occ < 1) { set_error("row kernel does not fit on an SM"); return NGSX_ERR_CUDA; }
  unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(A.nchunks, (int64_t)sm_count * occ));
  kern<<<grid, RT, smem, st>>>(A);
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
  RowArgs A{P->nchunks_active, P->chunk_order, P->chunk_q, P->chunk_split, P->rows, P->rowrec, reinterpret_cast<const uint4 *>(P->desc), P->rowptr, P->G, P->cutmats, P->tab,
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
  cudaFree(P->rows); cudaFree(P->pairptr); cudaFree(P->desc); cudaFree(P->rowrec); cudaFree(P->chunk_q); cudaFree(P->chunk_order); cudaFree(P->chunk_split);
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
    PLAN_CUDA(cudaMalloc(&P->chunk_split, P->nchunks * (RT + 1) * sizeof(int32_t)));
    k_chunk_split<<<(unsigned)((P->nchunks * (RT + 1) + 255) / 256), 256, 0, st>>>(P->nchunks, P->chunk_q, P->pairptr, P->chunk_split);
    // chunk processing order
    uint32_t *ck = nullptr, *ck2 = nullptr;
    int32_t *ci2 = nullptr;
    PLAN_CUDA(cudaMalloc(&ck, P->nchunks * sizeof(uint32_t)));
    PLAN_CUDA(cudaMalloc(&ck2, P->nchunks * sizeof(uint32_t)));
    PLAN_CUDA(cudaMalloc(&ci2, P->nchunks * sizeof(int32_t)));
    PLAN_CUDA(cudaMalloc(&P->chunk_order, P->nchunks * sizeof(int32_t)));
    PLAN_CUDA(cudaMemsetAsync(d_stat, 0, sizeof(unsigned long long), st));
    k_chunk_keys<<<(unsigned)((P->nchunks + 255) / 256), 256, 0, st>>>(P->nchunks, P->chunk_q, P->pairptr, P->desc, P->ds, ck, ci2, d_stat);
    cudaFree(tmp); tmp = nullptr; tmp_bytes = 0;
    PLAN_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, ck, ck2, ci2, P->chunk_order, (int)P->nchunks, 0, 32, st));
    PLAN_CUDA(cudaMalloc(&tmp, tmp_bytes));
    PLAN_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, ck, ck2, ci2, P->chunk_order, (int)P->nchunks, 0, 32, st));
    g_launches.fetch_add(5);
    unsigned long long nonempty = 0;
    PLAN_CUDA(cudaMemcpyAsync(&nonempty, d_stat, sizeof nonempty, cudaMemcpyDeviceToHost, st));
    PLAN_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof bad, cudaMemcpyDeviceToHost, st));
    PLAN_CUDA(cudaStreamSynchronize(st));
    cudaFree(ck); cudaFree(ck2); cudaFree(ci2);
    PLAN_CUDA(cudaGetLastError());
    P->nchunks_active = (int64_t)nonempty;
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
