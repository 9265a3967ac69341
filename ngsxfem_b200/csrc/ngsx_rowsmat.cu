// ngsx_rowsmat.cu -- owner-computes gather of PRECOMPUTED local matrices into CSR (the general form of d-3,
// SparseMatrixTM::AddElementMatrix, NGS/linalg/sparsematrix_impl.hpp:63-103): every listed entity k (an element, or a facet
// patch with its merged dof list, NGS/comp/bilinearform.cpp:1800-1853) owns a dense nd x nd block; the plan
// (ngsx_rowplan_create) schedules per CSR row the (entity, local row) pairs that contribute to it.  Same tile / lane / step
// structure as k_rows (ngsx_rows.cu): warp = tile of consecutive rows in a shared-memory window, lane = the rows the schedule
// dealt to it, pairs summed in plan order -> each CSR value written once per call, bitwise reproducible, no FP atomics.
//
// Two sources of the block rows:
//   * dense blocks mats[k][i][j] (ngsx_assemble_csr_rows_mats): a lane fetches its nd-double row from a lane-private address --
//     32 distinct lines per load instruction, and sector groups of 128 bytes for 80-byte rows (round 1: 3.0 GB read for 1.8 GB
//     of blocks, memory-instruction queue saturated);
//   * PACKED rows (ngsx_assemble_csr_rows_packed): the producer kernel writes row i of block k at slot * ndp, slot = the position
//     of the pair (k, i) in the schedule (ngsx_rowplan_rowslots).  The rows a warp needs in one step are then ONE contiguous run:
//     the warp copies it with coalesced 128-bit loads into its staging area and every lane reads its row back with
//     128-bit shared-memory loads; the next step's run is in flight (registers) during the window adds.
//     Every byte is read once, in whole lines.
#include "ngsx_rows.cuh"

namespace ngsx {

struct RowMatArgs {
  int64_t ntiles;
  const TileHdr *tiles;
  const uint8_t *desc;
  const double *mats;      // dense blocks, or packed rows
  double *vals;
  int accumulate;
};

// descriptor of DS bytes as DS/16 vector loads
template <int DS> struct Desc { uint4 w[DS / 16]; };

template <int DS>
__device__ __forceinline__ void ld_desc_if(bool p, const uint8_t *desc, int64_t slot, Desc<DS> &d) {
  const uint4 *q = reinterpret_cast<const uint4 *>(desc + slot * DS);
#pragma unroll
  for (int t = 0; t < DS / 16; t++) ldg128_if(p, q + t, d.w[t]);
}

template <int DS>
__device__ __forceinline__ Desc<DS> ld_desc(const uint8_t *desc, int64_t slot) {
  Desc<DS> d;
  const uint4 *q = reinterpret_cast<const uint4 *>(desc + slot * DS);
#pragma unroll
  for (int t = 0; t < DS / 16; t++) d.w[t] = __ldg(q + t);
  return d;
}

template <int DS, int B>
__device__ __forceinline__ uint32_t desc_byte(const Desc<DS> &d) {
  constexpr int wi = B >> 2, v = wi >> 2, c = wi & 3;
  const uint32_t x = c == 0 ? d.w[v].x : (c == 1 ? d.w[v].y : (c == 2 ? d.w[v].z : d.w[v].w));
  return __byte_perm(x, 0u, 0x4440u | (uint32_t)(B & 3));
}

template <int ND, int DS, int J>
__device__ __forceinline__ void desc_addr(const Desc<DS> &d, uint32_t row_addr, uint32_t *addr) {
  if constexpr (J < ND) { addr[J] = row_addr + 8u * desc_byte<DS, 6 + J>(d); desc_addr<ND, DS, J + 1>(d, row_addr, addr); }
}

__device__ __forceinline__ double lds_f64(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_f64(uint32_t addr, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v)); }

template <int ND>
__device__ __forceinline__ void ld_matrow(const double *m, double *v) {
  if constexpr (ND % 2 == 0) {
#pragma unroll
    for (int j = 0; j < ND; j += 2) {
      const double2 t = __ldg(reinterpret_cast<const double2 *>(m + j));
      v[j] = t.x; v[j + 1] = t.y;
    }
  } else {
#pragma unroll
    for (int j = 0; j < ND; j++) v[j] = __ldg(m + j);
  }
}

// window adds of one pair: the nd columns of one entity are distinct dofs -- read all, add, write all
template <int ND, int DS>
__device__ __forceinline__ void add_row(const Desc<DS> &d, uint32_t acc_addr, const double *v) {
  uint32_t addr[ND];
  desc_addr<ND, DS, 0>(d, acc_addr + 8u * (d.w[0].y & 0xffffu), addr);
  double a[ND];
#pragma unroll
  for (int j = 0; j < ND; j++) a[j] = lds_f64(addr[j]);
#pragma unroll
  for (int j = 0; j < ND; j++) sts_f64(addr[j], a[j] + v[j]);
}

// ---- dense blocks ---------------------------------------------------------------------------------------------------
template <int ND>
__global__ void __launch_bounds__(RW * 32) k_rows_mats(RowMatArgs A) {
  constexpr int DS = ND <= 10 ? 16 : 32;
  extern __shared__ __align__(16) double smem_rm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *acc = smem_rm + warp * ACCW;
  const uint32_t acc_addr = (uint32_t)__cvta_generic_to_shared(acc);
  const int64_t nwarps = (int64_t)gridDim.x * RW;
  for (int64_t tt = (int64_t)blockIdx.x * RW + warp; tt < A.ntiles; tt += nwarps) {
    const TileHdr *H = A.tiles + tt;
    const int64_t z0 = __ldg(&H->z0);
    const int zlen = __ldg(&H->zlen), nsteps = __ldg(&H->nsteps);
    const int myload = __ldg(&H->load[lane]);
    int64_t base = __ldg(&H->d0);
    prefetch_block_l2(A.desc, base, __reduce_add_sync(0xffffffffu, myload), DS, lane);
    if (tt + nwarps < A.ntiles) prefetch_l1(A.tiles + tt + nwarps);
    if (A.accumulate) for (int t = lane; t < zlen; t += 32) acc[t] = A.vals[z0 + t];
    else for (int t = lane; t < zlen; t += 32) acc[t] = 0.0;
    Desc<DS> d{};
    double v[ND];
    if (0 < myload) {
      d = ld_desc<DS>(A.desc, base + lane);
      ld_matrow<ND>(A.mats + (int64_t)(d.w[0].x & DESC_KMASK) * (ND * ND) + (int)(d.w[0].x >> DESC_KBITS) * ND, v);
    }
    base += __popc(__ballot_sync(0xffffffffu, 0 < myload));
    __syncwarp();
    for (int s = 0; s < nsteps; s++) {
      const Desc<DS> dc = d;
      double cur[ND];
#pragma unroll
      for (int j = 0; j < ND; j++) cur[j] = v[j];
      if (s + 1 < myload) {                                      // next pair's descriptor and matrix row fly during the adds
        d = ld_desc<DS>(A.desc, base + lane);
        ld_matrow<ND>(A.mats + (int64_t)(d.w[0].x & DESC_KMASK) * (ND * ND) + (int)(d.w[0].x >> DESC_KBITS) * ND, v);
      }
      base += __popc(__ballot_sync(0xffffffffu, s + 1 < myload));
      if (s < myload) add_row<ND, DS>(dc, acc_addr, cur);
    }
    __syncwarp();
    for (int t = lane; t < zlen; t += 32) A.vals[z0 + t] = acc[t];
    __syncwarp();
  }
}

// ---- packed rows ----------------------------------------------------------------------------------------------------
// Row stride NDP = nd rounded up to even: every step's run then starts on a 16-byte boundary.  Staging area of a warp: the
// rows of one step, 32 x NDP doubles = NDP/2 16-byte chunks per lane.
template <int ND> struct Packed {
  static constexpr int NDP = ND + (ND & 1);
  static constexpr int NC = NDP / 2;                              // 16-byte chunks per lane of a full step
  static constexpr int STAGE = 32 * NDP;                          // doubles
};

template <int ND>
__global__ void __launch_bounds__(RW * 32) k_rows_packed(RowMatArgs A) {
  constexpr int DS = ND <= 10 ? 16 : 32;
  constexpr int NDP = Packed<ND>::NDP, NC = Packed<ND>::NC, STAGE = Packed<ND>::STAGE;
  extern __shared__ __align__(16) double smem_rp[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *acc = smem_rp + warp * (ACCW + STAGE);
  double *stage = acc + ACCW;
  const uint32_t acc_addr = (uint32_t)__cvta_generic_to_shared(acc);
  const int64_t nwarps = (int64_t)gridDim.x * RW;
  for (int64_t tt = (int64_t)blockIdx.x * RW + warp; tt < A.ntiles; tt += nwarps) {
    const TileHdr *H = A.tiles + tt;
    const int64_t z0 = __ldg(&H->z0);
    const int zlen = __ldg(&H->zlen), nsteps = __ldg(&H->nsteps);
    const int myload = __ldg(&H->load[lane]);
    int64_t base = __ldg(&H->d0);                                 // slot of (step, lane 0)
    {                                                             // the tile's descriptors and rows are two contiguous blocks: into L2 now
      const int tot = __reduce_add_sync(0xffffffffu, myload);
      prefetch_block_l2(A.desc, base, tot, DS, lane);
      prefetch_block_l2(A.mats, base, tot, NDP * 8, lane);
      if (tt + nwarps < A.ntiles) prefetch_l1(A.tiles + tt + nwarps);
    }
    if (A.accumulate) for (int t = lane; t < zlen; t += 32) acc[t] = A.vals[z0 + t];
    else for (int t = lane; t < zlen; t += 32) acc[t] = 0.0;
    // the run of step s: cnt(s) rows of NDP doubles at A.mats + base * NDP, i.e. cnt * NC chunks of 16 bytes.
    // Registers: descriptors two steps ahead, the run one step ahead.
    Desc<DS> d{}, dn{};
    uint4 chunk[NC];
    int cnt = __popc(__ballot_sync(0xffffffffu, 0 < myload));
    int64_t dbase = base + cnt;                                   // descriptor slot of (step 1, lane 0)
    {
      ld_desc_if<DS>(0 < myload, A.desc, base + lane, d);
      ld_desc_if<DS>(1 < myload, A.desc, dbase + lane, dn);
      dbase += __popc(__ballot_sync(0xffffffffu, 1 < myload));
      const uint4 *src = reinterpret_cast<const uint4 *>(A.mats + base * NDP);
#pragma unroll
      for (int c = 0; c < NC; c++) if (lane + 32 * c < cnt * NC) chunk[c] = __ldg(src + lane + 32 * c);
    }
    __syncwarp();
    for (int s = 0; s < nsteps; s++) {
      const Desc<DS> dc = d;
      {                                                           // park the run of this step
        uint4 *dst = reinterpret_cast<uint4 *>(stage);
#pragma unroll
        for (int c = 0; c < NC; c++) if (lane + 32 * c < cnt * NC) dst[lane + 32 * c] = chunk[c];
      }
      base += cnt;
      cnt = __popc(__ballot_sync(0xffffffffu, s + 1 < myload));
      d = dn;
      ld_desc_if<DS>(s + 2 < myload, A.desc, dbase + lane, dn);
      dbase += __popc(__ballot_sync(0xffffffffu, s + 2 < myload));
      if (cnt > 0) {                                              // next step's run flies during the adds
        const uint4 *src = reinterpret_cast<const uint4 *>(A.mats + base * NDP);
#pragma unroll
        for (int c = 0; c < NC; c++) if (lane + 32 * c < cnt * NC) chunk[c] = __ldg(src + lane + 32 * c);
      }
      __syncwarp();
      if (s < myload) {
        double v[NDP];
        const double2 *row = reinterpret_cast<const double2 *>(stage + lane * NDP);
#pragma unroll
        for (int j = 0; j < NC; j++) { const double2 t = row[j]; v[2 * j] = t.x; v[2 * j + 1] = t.y; }
        add_row<ND, DS>(dc, acc_addr, v);
      }
      __syncwarp();
    }
    for (int t = lane; t < zlen; t += 32) A.vals[z0 + t] = acc[t];
    __syncwarp();
  }
}

template <int ND, bool PACKED>
static int launch_rows_mats_t(const RowMatArgs &A, int sm_count, cudaStream_t st) {
  const size_t smem = (size_t)RW * (ACCW + (PACKED ? Packed<ND>::STAGE : 0)) * sizeof(double);
  auto kern = PACKED ? k_rows_packed<ND> : k_rows_mats<ND>;
  NGSX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  NGSX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, RW * 32, smem));
  if (occ < 1) { set_error("matrix row gather does not fit on an SM"); return NGSX_ERR_CUDA; }
  const int64_t need = (A.ntiles + RW - 1) / RW;
  const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)sm_count * occ));
  kern<<<grid, RW * 32, smem, st>>>(A);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

// vectors: thread per dof, contributions summed in list order
__global__ void __launch_bounds__(256) k_gather_vec(int64_t ndof, const int64_t *__restrict__ dofptr, const int64_t *__restrict__ pairs,
                                                   const double *__restrict__ src, double scale, int accumulate, double *__restrict__ out) {
  const int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= ndof) return;
  const int64_t a = dofptr[d], b = dofptr[d + 1];
  if (a == b && accumulate) return;
  double s = 0.0;
  for (int64_t p = a; p < b; p++) s += __ldg(src + __ldg(pairs + p));
  out[d] = accumulate ? out[d] + scale * s : scale * s;
}

template <bool PACKED>
static int dispatch_rows_mats(ngsx_rowplan *plan, const double *mats, int accumulate, double *vals, void *stream) {
  if (plan->nrows == 0 || plan->nlist == 0) return NGSX_OK;
  NGSX_REQUIRE(mats && vals, "assemble_csr_rows_mats: mats / vals are NULL");
  if (plan->lower) { set_error("assemble_csr_rows_mats: lower-triangle plans are served by ngsx_assemble_csr_rows_h1 only"); return NGSX_ERR_UNSUPPORTED; }
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  RowMatArgs A{plan->ntiles, plan->tiles, plan->desc, mats, vals, accumulate};
  cudaStream_t st = as_stream(stream);
  switch (plan->nd) {
    case 3: return launch_rows_mats_t<3, PACKED>(A, dd->sm_count, st);
    case 4: return launch_rows_mats_t<4, PACKED>(A, dd->sm_count, st);
    case 5: return launch_rows_mats_t<5, PACKED>(A, dd->sm_count, st);
    case 6: return launch_rows_mats_t<6, PACKED>(A, dd->sm_count, st);
    case 8: return launch_rows_mats_t<8, PACKED>(A, dd->sm_count, st);
    case 9: return launch_rows_mats_t<9, PACKED>(A, dd->sm_count, st);
    case 10: return launch_rows_mats_t<10, PACKED>(A, dd->sm_count, st);
    case 12: return launch_rows_mats_t<12, PACKED>(A, dd->sm_count, st);
    case 14: return launch_rows_mats_t<14, PACKED>(A, dd->sm_count, st);
    case 16: return launch_rows_mats_t<16, PACKED>(A, dd->sm_count, st);
    case 18: return launch_rows_mats_t<18, PACKED>(A, dd->sm_count, st);
    case 20: return launch_rows_mats_t<20, PACKED>(A, dd->sm_count, st);
  }
  set_error("assemble_csr_rows_mats: unsupported block size (3, 4, 5, 6, 8, 9, 10, 12, 14, 16, 18, 20)");
  return NGSX_ERR_UNSUPPORTED;
}

}  // namespace ngsx

using namespace ngsx;

extern "C" int ngsx_gather_vec(int64_t ndof, const int64_t *dofptr, const int64_t *pairs, const double *src, double scale, int accumulate,
                               double *out, void *stream) {
  NGSX_REQUIRE(ndof >= 0 && (ndof == 0 || (dofptr && out)), "gather_vec: bad arguments");
  if (ndof == 0) return NGSX_OK;
  k_gather_vec<<<(unsigned)((ndof + 255) / 256), 256, 0, as_stream(stream)>>>(ndof, dofptr, pairs, src, scale, accumulate, out);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

extern "C" int ngsx_assemble_csr_rows_mats(ngsx_rowplan *plan, const double *mats, int accumulate, double *vals, void *stream) {
  NGSX_REQUIRE(plan, "assemble_csr_rows_mats: plan is NULL");
  return dispatch_rows_mats<false>(plan, mats, accumulate, vals, stream);
}

extern "C" int ngsx_assemble_csr_rows_packed(ngsx_rowplan *plan, const double *rows, int accumulate, double *vals, void *stream) {
  NGSX_REQUIRE(plan, "assemble_csr_rows_packed: plan is NULL");
  NGSX_REQUIRE((reinterpret_cast<uintptr_t>(rows) & 15) == 0, "assemble_csr_rows_packed: rows must be 16-byte aligned");
  return dispatch_rows_mats<true>(plan, rows, accumulate, vals, stream);
}
