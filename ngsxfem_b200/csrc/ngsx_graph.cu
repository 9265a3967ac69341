// ngsx_graph.cu -- (d-1) CSR sparsity of the marked elements / facet patches, built on the device:
// one 64-bit key (row<<32 | col) per coupling, radix sort, unique, row pointers by binary search.
// Replaces the TableCreator + MatrixGraph of RestrictedBilinearForm::GetGraph (utils/restrictedblf.cpp:22-138).
// CUB (part of the CUDA toolkit) provides the sort/unique primitives; this is setup work, not the timed hot loop.
#include <cub/cub.cuh>
#include <vector>
#include "ngsx_host.cuh"

struct ngsx_graph {
  int64_t ndof = 0;
  std::vector<std::pair<uint64_t *, int64_t>> chunks;   // device key buffers
  uint64_t *keys = nullptr;                             // sorted unique keys after finalize
  int64_t nnz = -1;
  int lower = 0;                                        // keep the couplings col <= row only (symmetric storage)
};

namespace ngsx {

__global__ void __launch_bounds__(256) k_make_keys(int64_t n, int nd, const int32_t *__restrict__ table, uint64_t *__restrict__ keys, int lower) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int ne = nd * nd;
  if (t >= n * ne) return;
  int64_t k = t / ne;
  int idx = (int)(t % ne);
  int32_t r = table[k * nd + idx / nd], c = table[k * nd + idx % nd];
  // negative dof numbers (unused / padded entries) are parked on an out-of-range key and dropped at finalize
  keys[t] = (r < 0 || c < 0 || (lower && c > r)) ? ~0ULL : (((uint64_t)(uint32_t)r << 32) | (uint32_t)c);
}

__global__ void __launch_bounds__(256) k_export_cols(int64_t nnz, const uint64_t *__restrict__ keys, int32_t *__restrict__ colidx) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nnz) colidx[t] = (int32_t)(keys[t] & 0xffffffffULL);
}

__global__ void __launch_bounds__(256) k_export_rowptr(int64_t ndof, int64_t nnz, const uint64_t *__restrict__ keys, int64_t *__restrict__ rowptr) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r > ndof) return;
  uint64_t target = (uint64_t)r << 32;
  int64_t lo = 0, hi = nnz;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (keys[mid] < target) lo = mid + 1; else hi = mid;
  }
  rowptr[r] = lo;
}

}  // namespace ngsx

using namespace ngsx;

extern "C" {

int ngsx_csr_graph_create(int64_t ndof, ngsx_graph **out) {
  NGSX_REQUIRE(out && ndof >= 0 && ndof < (1LL << 31), "csr_graph_create: bad arguments");
  *out = new ngsx_graph();
  (*out)->ndof = ndof;
  return NGSX_OK;
}

int ngsx_csr_graph_set_lower(ngsx_graph *g, int on) {
  NGSX_REQUIRE(g && g->chunks.empty() && g->nnz < 0, "csr_graph_set_lower: call before the first add_table");
  g->lower = on ? 1 : 0;
  return NGSX_OK;
}
int ngsx_csr_graph_add_table(ngsx_graph *g, int64_t n, int32_t nd, const int32_t *table, void *stream) {
  NGSX_REQUIRE(g && n >= 0 && nd > 0 && (n == 0 || table), "csr_graph_add_table: bad arguments");
  if (n == 0) return NGSX_OK;
  int64_t cnt = n * nd * nd;
  uint64_t *keys = nullptr;
  NGSX_CUDA(cudaMalloc(&keys, cnt * sizeof(uint64_t)));
  k_make_keys<<<(unsigned)((cnt + 255) / 256), 256, 0, as_stream(stream)>>>(n, nd, table, keys, g->lower);
  g->chunks.emplace_back(keys, cnt);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_csr_graph_finalize(ngsx_graph *g, int64_t *h_nnz, void *stream) {
  NGSX_REQUIRE(g && h_nnz, "csr_graph_finalize: bad arguments");
  cudaStream_t st = as_stream(stream);
  int64_t total = 0;
  for (auto &c : g->chunks) total += c.second;
  if (total == 0) { g->nnz = 0; *h_nnz = 0; return NGSX_OK; }
  uint64_t *in = nullptr, *alt = nullptr;
  if (g->chunks.size() == 1) {
    in = g->chunks[0].first;
  } else {
    NGSX_CUDA(cudaMalloc(&in, total * sizeof(uint64_t)));
    int64_t off = 0;
    for (auto &c : g->chunks) {
      NGSX_CUDA(cudaMemcpyAsync(in + off, c.first, c.second * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
      off += c.second;
    }
    NGSX_CUDA(cudaStreamSynchronize(st));
    for (auto &c : g->chunks) cudaFree(c.first);
  }
  g->chunks.clear();
  NGSX_CUDA(cudaMalloc(&alt, total * sizeof(uint64_t)));
  cub::DoubleBuffer<uint64_t> db(in, alt);
  int bits = 33;
  while (bits < 64 && (1LL << (bits - 32)) < g->ndof + 1) bits++;
  size_t tmp_bytes = 0;
  NGSX_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, db, total, 0, 64, st));
  void *tmp = nullptr;
  NGSX_CUDA(cudaMalloc(&tmp, tmp_bytes));
  // padded keys are ~0: sort on all 64 bits only if padding can occur; (row,col) keys need `bits`
  NGSX_CUDA(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, db, total, 0, 64, st));
  cudaFree(tmp);
  uint64_t *sorted = db.Current(), *other = db.Alternate();
  int64_t *d_num = nullptr;
  NGSX_CUDA(cudaMalloc(&d_num, sizeof(int64_t)));
  tmp_bytes = 0;
  NGSX_CUDA(cub::DeviceSelect::Unique(nullptr, tmp_bytes, sorted, other, d_num, total, st));
  NGSX_CUDA(cudaMalloc(&tmp, tmp_bytes));
  NGSX_CUDA(cub::DeviceSelect::Unique(tmp, tmp_bytes, sorted, other, d_num, total, st));
  g_launches.fetch_add(8);
  int64_t nuniq = 0;
  NGSX_CUDA(cudaMemcpyAsync(&nuniq, d_num, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  NGSX_CUDA(cudaStreamSynchronize(st));
  cudaFree(tmp);
  cudaFree(d_num);
  cudaFree(sorted);
  // drop the parked padding key (largest, hence last) if present
  if (nuniq > 0) {
    uint64_t last = 0;
    NGSX_CUDA(cudaMemcpy(&last, other + (nuniq - 1), sizeof(uint64_t), cudaMemcpyDeviceToHost));
    if (last == ~0ULL) nuniq--;
  }
  g->keys = other;
  g->nnz = nuniq;
  *h_nnz = nuniq;
  return NGSX_OK;
}

int ngsx_csr_graph_export(ngsx_graph *g, int64_t *rowptr, int32_t *colidx, void *stream) {
  NGSX_REQUIRE(g && g->nnz >= 0 && rowptr && (g->nnz == 0 || colidx), "csr_graph_export: finalize first");
  cudaStream_t st = as_stream(stream);
  if (g->nnz > 0) {
    k_export_cols<<<(unsigned)((g->nnz + 255) / 256), 256, 0, st>>>(g->nnz, g->keys, colidx);
    NGSX_LAUNCH_CHECK();
  }
  k_export_rowptr<<<(unsigned)((g->ndof + 1 + 255) / 256), 256, 0, st>>>(g->ndof, g->nnz, g->keys, rowptr);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_csr_graph_destroy(ngsx_graph *g) {
  if (!g) return NGSX_OK;
  cudaDeviceSynchronize();
  for (auto &c : g->chunks) cudaFree(c.first);
  if (g->keys) cudaFree(g->keys);
  delete g;
  return NGSX_OK;
}

}  // extern "C"
