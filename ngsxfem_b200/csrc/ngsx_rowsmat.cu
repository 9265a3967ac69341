// ngsx_rowsmat.cu -- owner-computes gather of PRECOMPUTED local matrices into CSR (the general form of d-3,
// SparseMatrixTM::AddElementMatrix, NGS/linalg/sparsematrix_impl.hpp:63-103): every listed entity k (an element, or a facet
// patch with its merged dof list, NGS/comp/bilinearform.cpp:1800-1853) owns a dense nd x nd block mats[k]; the plan
// (ngsx_rowplan_create) lists per CSR row the (entity, local row) pairs that contribute to it.  Same tile / lane structure as
// k_rows (ngsx_rows.cu): warp = tile of consecutive rows in a shared-memory window, lane = contiguous row range, pairs summed
// in plan order -> each CSR value written once per call, bitwise reproducible, no floating-point atomics.
// Used for the ghost-penalty facet-patch operator (f-1), whose blocks come from k_facetpatch_c (ngsx_facetpatch.cu).
#include "ngsx_rows.cuh"

namespace ngsx {

struct RowMatArgs {
  int64_t ntiles;
  const TileInfo *info;
  const int32_t *split;
  const RowRec *rowrec;
  const uint8_t *desc;
  const double *mats;
  double *vals;
  int accumulate;
};

// descriptor of DS bytes as DS/16 vector loads
template <int DS> struct Desc { uint4 w[DS / 16]; };

template <int DS>
__device__ __forceinline__ Desc<DS> ld_desc(const uint8_t *desc, uint32_t p) {
  Desc<DS> d;
  const uint4 *q = reinterpret_cast<const uint4 *>(desc + (size_t)p * DS);
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
  if constexpr (J < ND) { addr[J] = row_addr + 8u * desc_byte<DS, 5 + J>(d); desc_addr<ND, DS, J + 1>(d, row_addr, addr); }
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

template <int ND>
__global__ void __launch_bounds__(RW * 32) k_rows_mats(RowMatArgs A) {
  constexpr int DS = ND <= 11 ? 16 : 32;
  extern __shared__ __align__(16) double smem_rm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *acc = smem_rm + warp * ACCW;
  const uint32_t acc_addr = (uint32_t)__cvta_generic_to_shared(acc);
  const int64_t nwarps = (int64_t)gridDim.x * RW;
  for (int64_t tt = (int64_t)blockIdx.x * RW + warp; tt < A.ntiles; tt += nwarps) {
    const TileInfo ti = A.info[tt];
    const int64_t z0 = ti.z0;
    const int zlen = ti.zlen;
    if (A.accumulate) for (int t = lane; t < zlen; t += 32) acc[t] = A.vals[z0 + t];
    else for (int t = lane; t < zlen; t += 32) acc[t] = 0.0;
    const int32_t *split = A.split + tt * 33;
    int q = __ldg(split + lane);
    const int qb = __ldg(split + lane + 1);
    __syncwarp();
    if (q < qb) {
      RowRec rec = ld_rowrec(A.rowrec + q);
      const RowRec recl = ld_rowrec(A.rowrec + qb - 1);
      uint32_t p = rec.p0, pend = rec.p0 + rec.np;
      const uint32_t plast = recl.p0 + recl.np;
      Desc<DS> d = ld_desc<DS>(A.desc, p);
      double v[ND];
      ld_matrow<ND>(A.mats + (int64_t)(int)d.w[0].x * (ND * ND) + (int)(d.w[0].y & 0xffu) * ND, v);
      while (p < plast) {
        if (p == pend) { q++; rec = ld_rowrec(A.rowrec + q); pend = rec.p0 + rec.np; }
        uint32_t addr[ND];
        desc_addr<ND, DS, 0>(d, acc_addr + 8u * (uint32_t)rec.off, addr);
        double cur[ND];
#pragma unroll
        for (int j = 0; j < ND; j++) cur[j] = v[j];
        if (p + 1 < plast) {                                     // next pair's descriptor and matrix row fly during the adds
          d = ld_desc<DS>(A.desc, p + 1);
          ld_matrow<ND>(A.mats + (int64_t)(int)d.w[0].x * (ND * ND) + (int)(d.w[0].y & 0xffu) * ND, v);
        }
        // the nd columns of one entity are distinct dofs: read all, add, write all
        double a[ND];
#pragma unroll
        for (int j = 0; j < ND; j++) a[j] = lds_f64(addr[j]);
#pragma unroll
        for (int j = 0; j < ND; j++) sts_f64(addr[j], a[j] + cur[j]);
        p++;
      }
    }
    __syncwarp();
    for (int t = lane; t < zlen; t += 32) A.vals[z0 + t] = acc[t];
    __syncwarp();
  }
}

template <int ND>
static int launch_rows_mats_t(const RowMatArgs &A, int sm_count, cudaStream_t st) {
  const size_t smem = (size_t)RW * ACCW * sizeof(double);
  auto kern = k_rows_mats<ND>;
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
  if (plan->nrows == 0 || plan->nlist == 0) return NGSX_OK;
  NGSX_REQUIRE(mats && vals, "assemble_csr_rows_mats: mats / vals are NULL");
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  RowMatArgs A{plan->ntiles_active, plan->tile_info, plan->tile_split, plan->rowrec, plan->desc, mats, vals, accumulate};
  cudaStream_t st = as_stream(stream);
  switch (plan->nd) {
    case 3: return launch_rows_mats_t<3>(A, dd->sm_count, st);
    case 4: return launch_rows_mats_t<4>(A, dd->sm_count, st);
    case 5: return launch_rows_mats_t<5>(A, dd->sm_count, st);
    case 6: return launch_rows_mats_t<6>(A, dd->sm_count, st);
    case 8: return launch_rows_mats_t<8>(A, dd->sm_count, st);
    case 9: return launch_rows_mats_t<9>(A, dd->sm_count, st);
    case 10: return launch_rows_mats_t<10>(A, dd->sm_count, st);
    case 12: return launch_rows_mats_t<12>(A, dd->sm_count, st);
    case 14: return launch_rows_mats_t<14>(A, dd->sm_count, st);
    case 20: return launch_rows_mats_t<20>(A, dd->sm_count, st);
  }
  set_error("assemble_csr_rows_mats: unsupported block size (3, 4, 5, 6, 8, 9, 10, 12, 14, 20)");
  return NGSX_ERR_UNSUPPORTED;
}
