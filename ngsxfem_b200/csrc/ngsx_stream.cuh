// ngsx_stream.cuh -- element-table streamer for the whole-mesh sweeps (classification, Integrate).
//
// These sweeps read the element -> vertex table once (16 B per tet: the only DRAM stream of any size, the per-vertex gathers
// hit L2) and were bound by the latency of the chain  table load -> level-set gather -> coordinate gather  at ~50 % occupancy
// (ncu: 21 % of DRAM peak, 12 warps stalled on the long scoreboard per issue).  Here a persistent CTA pulls its chunks of the
// table into shared memory with the bulk-copy engine (cp.async.bulk, completion on an mbarrier), NST chunks ahead of the
// threads, so the table stream never waits for a thread and a thread starts each chunk with its indices already on chip.
// A chunk is CH = 1024 consecutive elements = 32 words of every per-element bit array; warp w of the CTA owns words
// 4 w .. 4 w + 3 of the chunk, lane = bit, and issues the gathers of its four elements together.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/ngsx_abi.h"

namespace ngsx {

constexpr int STREAM_CH = 1024;      // elements per chunk
#ifndef NGSX_STREAM_NST
#define NGSX_STREAM_NST 4
#endif
constexpr int STREAM_NST = NGSX_STREAM_NST;   // chunks in flight per CTA
constexpr int STREAM_THREADS = 256;  // 8 warps x 4 words x 32 lanes = one chunk
constexpr int STREAM_WPW = 4;        // words per warp and chunk

template <int DIM> constexpr size_t stream_smem() { return (size_t)STREAM_NST * STREAM_CH * (DIM + 1) * 4 + STREAM_NST * 8; }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "NGSX_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra NGSX_DONE;\n"
      "bra NGSX_WAIT;\n"
      "NGSX_DONE:\n"
      "}\n" ::"r"(bar), "r"(parity)
      : "memory");
}
// global -> shared bulk copy (bytes: multiple of 16, both addresses 16-byte aligned), completion counted on the mbarrier
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// vertex numbers of element i of the chunk: from the staged bytes, or (last elements of a 2D tail chunk whose 12-byte record
// crosses the 16-byte granule of the bulk copy) straight from the table
template <int DIM>
__device__ __forceinline__ void stream_vertices(const int32_t *stage, int nstaged, const int32_t *el2v, int64_t first, int i, int *v) {
  if constexpr (DIM == 3) {
    const int4 q = i < nstaged ? reinterpret_cast<const int4 *>(stage)[i] : __ldg(reinterpret_cast<const int4 *>(el2v) + first + i);
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  } else {
    const int32_t *q = i < nstaged ? stage + 3 * i : el2v + 3 * (first + i);
    v[0] = q[0]; v[1] = q[1]; v[2] = q[2];
  }
}

// body(c, first, n, stage, nstaged): chunk c holds elements [first, first + n); called by all threads of the CTA
template <int DIM, typename F>
__device__ __forceinline__ void stream_elements(const ngsx_mesh &M, unsigned char *smem, F &&body) {
  constexpr int ELB = (DIM + 1) * 4, STAGE = STREAM_CH * ELB;
  const uint32_t s_base = (uint32_t)__cvta_generic_to_shared(smem);
  const uint32_t s_bar = s_base + STREAM_NST * STAGE;
  const int64_t nchunks = (M.ne + STREAM_CH - 1) / STREAM_CH;
  const unsigned char *g = reinterpret_cast<const unsigned char *>(M.el2v);
  // a 2D table that does not start on a 16-byte boundary (element blocks of a partition) is read directly, nothing is staged
  const uint32_t granule = (reinterpret_cast<uintptr_t>(g) & 15u) == 0 ? ~15u : 0u;
  auto issue = [&](int64_t c, int st) {
    const int64_t first = c * STREAM_CH;
    const int64_t n = M.ne - first < STREAM_CH ? M.ne - first : STREAM_CH;
    const uint32_t bytes = (uint32_t)(n * ELB) & granule;
    if (bytes) {
      mbar_expect_tx(s_bar + 8 * st, bytes);
      bulk_g2s(s_base + st * STAGE, g + first * ELB, bytes, s_bar + 8 * st);
    } else {
      mbar_arrive(s_bar + 8 * st);
    }
  };
  if (threadIdx.x == 0) {
    for (int st = 0; st < STREAM_NST; st++) mbar_init(s_bar + 8 * st, 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int st = 0; st < STREAM_NST; st++) {
      const int64_t c = blockIdx.x + (int64_t)st * gridDim.x;
      if (c < nchunks) issue(c, st);
    }
  }
  int it = 0;
  for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x, it++) {
    const int st = it % STREAM_NST;
    mbar_wait(s_bar + 8 * st, (uint32_t)(it / STREAM_NST) & 1u);
    const int64_t first = c * STREAM_CH;
    const int n = (int)(M.ne - first < STREAM_CH ? M.ne - first : STREAM_CH);
    body(c, first, n, reinterpret_cast<const int32_t *>(smem + (size_t)st * STAGE), (int)(((uint32_t)(n * ELB) & granule) / ELB));
    __syncthreads();                                  // every thread is done with the stage: refill it
    if (threadIdx.x == 0) {
      const int64_t c2 = c + (int64_t)STREAM_NST * gridDim.x;
      if (c2 < nchunks) issue(c2, st);
    }
  }
}

// set bits of a chunk of 32 mask words, dealt out 32 at a time in (word, bit) order: item t of the chunk -> (word, bit).
// cw: this lane's word; end/begin: inclusive/exclusive prefix sums of the bit counts over the lanes.
struct BitDeal {
  uint32_t cw;
  int cnt, begin, end, total;
  __device__ __forceinline__ void scan(uint32_t word, int lane) {
    cw = word;
    cnt = __popc(cw);
    end = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, end, o); if (lane >= o) end += t; }
    total = __shfl_sync(0xffffffffu, end, 31);
    begin = end - cnt;
  }
  // all lanes call; returns the bit (0..31) of item t and its word in `owner`, or -1 when t >= total
  __device__ __forceinline__ int item(int t, int &owner) const {
    owner = 0;
#pragma unroll
    for (int step = 16; step > 0; step >>= 1) {
      const int probe = owner + step - 1;
      const int pe = __shfl_sync(0xffffffffu, end, probe & 31);
      if (probe < 32 && pe <= t) owner += step;
    }
    owner = owner < 31 ? owner : 31;
    const uint32_t ow = __shfl_sync(0xffffffffu, cw, owner);
    const int ob = __shfl_sync(0xffffffffu, begin, owner);
    return t < total ? (int)__fns(ow, 0u, t - ob + 1) : -1;
  }
};

}  // namespace ngsx
