// ngsx_peer.cu -- the path's only exchange between GPUs: the scalar results of Integrate() (a few doubles per step and rank),
// summed over the ranks of one node WITHOUT a collective library call.
//
// With NCCL the 16-byte all-reduce costs more than its bytes: a host-side enqueue per step and rank, a resident kernel that
// spins for the slowest peer beside the rank's persistent gather -- 0.173 -> 0.214 ms per step at 8 GPUs.  Here the reduction
// kernel's output is PUBLISHED into every peer's mailbox with plain stores over NVLink (CUDA IPC mappings of the peers' device
// buffers), as the last node of the rank's captured step; the sums of the PREVIOUS step are collected in the same launch, when
// they have long arrived, so no rank ever waits for another inside a step.  Summation in rank order: every rank computes
// bitwise the same totals, run to run.
//
// Mailbox of a rank: [NGSX_PEER_SLOTS step slots][world ranks]{ double v[NGSX_PEER_MAXVALS]; int64 tag }.  A writer stores the
// values, fences system-wide, then stores tag = step; a reader polls the tag.  Four slots: a rank can be at most two steps ahead
// of the slowest one before its own collect of the previous step holds it (see bench.py), so a slot is never overwritten unread.
#include <cstring>
#include "ngsx_host.cuh"

namespace ngsx {

constexpr int PEER_SLOTS = 4;
struct PeerCell { double v[NGSX_PEER_MAXVALS]; long long tag; };

__device__ __forceinline__ PeerCell *peer_cell(void *mailbox, int world, long long step, int rank) {
  return reinterpret_cast<PeerCell *>(mailbox) + (step & (PEER_SLOTS - 1)) * world + rank;
}

// sums of step `step` from this rank's own mailbox (all `world` cells tagged with it), rank order
__device__ __forceinline__ void peer_collect(void *mailbox, int n, int world, long long step, double *out) {
  __shared__ int ok;
  const int t = threadIdx.x;
  if (t == 0) ok = 1;
  __syncthreads();
  if (t < world) {
    volatile long long *tag = &peer_cell(mailbox, world, step, t)->tag;
    long long spins = 0;
    while (*tag != step && ++spins < (1LL << 28)) {}
    if (*tag != step) ok = 0;                                       // a peer never published: poison the result instead of hanging
  }
  __syncthreads();
  __threadfence_system();
  if (t < n) {
    double s = 0.0;
    for (int r = 0; r < world; r++) s += *reinterpret_cast<volatile double *>(&peer_cell(mailbox, world, step, r)->v[t]);
    out[t] = ok ? s : __longlong_as_double(0x7ff8000000000000LL);
  }
}

__global__ void __launch_bounds__(64) k_peer_exchange(const double *__restrict__ local, int n, int rank, int world, void *const *__restrict__ peers,
                                                     void *mailbox, long long *d_step, double *out_prev) {
  const int t = threadIdx.x;
  const long long step = *d_step + 1;
  if (step > 1) peer_collect(mailbox, n, world, step - 1, out_prev);
  if (t < world) {                                                  // publish: thread t serves peer t (its own mailbox included)
    PeerCell *c = peer_cell(peers[t], world, step, rank);
    for (int i = 0; i < n; i++) *reinterpret_cast<volatile double *>(&c->v[i]) = local[i];
    __threadfence_system();
    *reinterpret_cast<volatile long long *>(&c->tag) = step;
  }
  __syncthreads();
  if (t == 0) *d_step = step;
}

__global__ void __launch_bounds__(64) k_peer_collect(void *mailbox, int n, int world, const long long *d_step, double *out) {
  peer_collect(mailbox, n, world, *d_step, out);
}

}  // namespace ngsx

using namespace ngsx;

extern "C" {

int ngsx_peer_mailbox_bytes(int world) { return world > 0 ? (int)(PEER_SLOTS * world * sizeof(PeerCell)) : 0; }

int ngsx_peer_alloc(int world, void **mailbox, unsigned char *handle64) {
  NGSX_REQUIRE(world > 0 && world <= 64 && mailbox && handle64, "peer_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void *p = nullptr;
  NGSX_CUDA(cudaMalloc(&p, (size_t)ngsx_peer_mailbox_bytes(world)));
  NGSX_CUDA(cudaMemset(p, 0, (size_t)ngsx_peer_mailbox_bytes(world)));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) { cudaFree(p); set_error(std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e)); return NGSX_ERR_CUDA; }
  std::memcpy(handle64, &h, 64);
  *mailbox = p;
  return NGSX_OK;
}

int ngsx_peer_open(const unsigned char *handle64, void **mailbox) {
  NGSX_REQUIRE(handle64 && mailbox, "peer_open: bad arguments");
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle64, 64);
  NGSX_CUDA(cudaIpcOpenMemHandle(mailbox, h, cudaIpcMemLazyEnablePeerAccess));
  return NGSX_OK;
}

int ngsx_peer_close(void *mailbox) {
  if (mailbox) NGSX_CUDA(cudaIpcCloseMemHandle(mailbox));
  return NGSX_OK;
}

int ngsx_peer_free(void *mailbox) {
  if (mailbox) NGSX_CUDA(cudaFree(mailbox));
  return NGSX_OK;
}

int ngsx_peer_exchange(const double *local, int n, int rank, int world, void *const *d_peers, void *mailbox, int64_t *d_step, double *out_prev,
                       void *stream) {
  NGSX_REQUIRE(local && d_peers && mailbox && d_step && out_prev && n > 0 && n <= NGSX_PEER_MAXVALS && world > 0 && world <= 64 &&
               rank >= 0 && rank < world, "peer_exchange: bad arguments");
  k_peer_exchange<<<1, 64, 0, as_stream(stream)>>>(local, n, rank, world, d_peers, mailbox, reinterpret_cast<long long *>(d_step), out_prev);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_peer_collect(void *mailbox, int n, int world, const int64_t *d_step, double *out, void *stream) {
  NGSX_REQUIRE(mailbox && d_step && out && n > 0 && n <= NGSX_PEER_MAXVALS && world > 0 && world <= 64, "peer_collect: bad arguments");
  k_peer_collect<<<1, 64, 0, as_stream(stream)>>>(mailbox, n, world, reinterpret_cast<const long long *>(d_step), out);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

}  // extern "C"
