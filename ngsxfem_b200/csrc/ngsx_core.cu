// ngsx_core.cu -- library plumbing, InterpolateToP1, CutInfo classification, straight-cut rule emission, Integrate.
// Hand-written CUDA for sm_100a; HBM-bound integer/FP64 streaming kernels (one thread per element, warp ballots
// for bit arrays, one warp per cut simplex for rule emission).
#include <cub/cub.cuh>
#include <algorithm>
#include <cstring>
#include <mutex>
#include <vector>
#include "ngsx_host.cuh"
#include "ngsx_stream.cuh"
#ifndef NGSX_INT_MINB
#define NGSX_INT_MINB 2      // resident CTAs per SM of the streamed sweeps (64 KB of stages each; 3 forces 80 registers and spills)
#endif
#ifndef NGSX_CLS_MINB
#define NGSX_CLS_MINB 2
#endif
#include "rule_tables.h"

namespace ngsx {

// =============================================================================================================
// host plumbing
// =============================================================================================================
static thread_local std::string t_err;
std::atomic<int64_t> g_launches{0};
void set_error(const std::string &msg) { t_err = msg; }

// ---- decomposition table: replay of LevelsetCutSimplex::Decompose (cutint/straightcutrule.cpp:131-196) on symbols
static void build_cut_entry(int dim, int dt, int mask, CutEntry &E) {
  memset(&E, 0, sizeof E);
  int nv = dim + 1;
  bool pos[4] = {false, false, false, false};
  for (int k = 0; k < nv; k++) pos[k] = (mask >> k) & 1;
  uint8_t cuts[6];
  int nc = 0;
  for (int i = 0; i < nv; i++)                       // CalcIFPolytopEUsingLset: i<j loop, ">= 0" differs (:53-60)
    for (int j = i + 1; j < nv; j++)
      if (pos[i] != pos[j]) cuts[nc++] = uint8_t((i << 4) | j);
  if (nc == 0) return;
  auto V = [](int k) { return uint8_t((k << 4) | k); };
  if (dt == IF) {
    E.npt = uint8_t(dim);
    if (nc == dim) {                                 // :139
      E.ns = 1;
      for (int k = 0; k < nc; k++) E.v[0][k] = cuts[k];
    } else if (nc == 4 && dim == 3) {                // :140-143
      E.ns = 2;
      E.v[0][0] = cuts[0]; E.v[0][1] = cuts[1]; E.v[0][2] = cuts[3];
      E.v[1][0] = cuts[0]; E.v[1][1] = cuts[2]; E.v[1][2] = cuts[3];
    }
    return;
  }
  E.npt = uint8_t(dim + 1);
  int rel[4], nrel = 0;
  for (int k = 0; k < nv; k++)
    if ((dt == POS && pos[k]) || (dt == NEG && !pos[k])) rel[nrel++] = k;
  if (nrel == 1) {                                   // :156-160  {cuts..., v}
    E.ns = 1;
    for (int k = 0; k < nc; k++) E.v[0][k] = cuts[k];
    E.v[0][nc] = V(rel[0]);
  } else if (nrel == 2 && dim == 2) {                // :161-168
    E.ns = 2;
    E.v[0][0] = V(rel[0]); E.v[0][1] = V(rel[1]); E.v[0][2] = cuts[1];
    E.v[1][0] = cuts[0];   E.v[1][1] = cuts[1];   E.v[1][2] = V(rel[0]);
  } else if (nrel == 2 && dim == 3) {                // :169-178
    E.ns = 3;
    E.v[0][0] = V(rel[1]); E.v[0][1] = cuts[1];   E.v[0][2] = cuts[2]; E.v[0][3] = cuts[3];
    E.v[1][0] = V(rel[0]); E.v[1][1] = V(rel[1]); E.v[1][2] = cuts[1]; E.v[1][3] = cuts[2];
    E.v[2][0] = cuts[0];   E.v[2][1] = cuts[1];   E.v[2][2] = cuts[2]; E.v[2][3] = V(rel[0]);
  } else if (nrel == 3 && dim == 3) {                // :179-188
    E.ns = 3;
    E.v[0][0] = cuts[0];   E.v[0][1] = cuts[1];   E.v[0][2] = cuts[2];   E.v[0][3] = V(rel[2]);
    E.v[1][0] = V(rel[0]); E.v[1][1] = V(rel[1]); E.v[1][2] = V(rel[2]); E.v[1][3] = cuts[1];
    E.v[2][0] = cuts[0];   E.v[2][1] = cuts[1];   E.v[2][2] = V(rel[0]); E.v[2][3] = V(rel[2]);
  }
}

static std::mutex g_dd_mutex;
static DeviceData g_dd[64];
static bool g_dd_ready[64];
static int g_rule_base[4];     // start of the rules of simplex dimension 0..3 inside DeviceData::rules

const DeviceData *device_data() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("cudaGetDevice failed"); return nullptr; }
  std::lock_guard<std::mutex> lock(g_dd_mutex);
  if (g_dd_ready[dev]) return &g_dd[dev];
  std::vector<double4> all;
  g_rule_base[0] = 0;
  all.push_back(make_double4(0, 0, 0, 1.0));                        // ET_POINT rule
  auto push = [&](const double (*d)[4], int n) { for (int i = 0; i < n; i++) all.push_back(make_double4(d[i][0], d[i][1], d[i][2], d[i][3])); };
  g_rule_base[1] = (int)all.size(); push(NGSX_RULE_SEGM_DATA, NGSX_RULE_SEGM_NPTS);
  g_rule_base[2] = (int)all.size(); push(NGSX_RULE_TRIG_DATA, NGSX_RULE_TRIG_NPTS);
  g_rule_base[3] = (int)all.size(); push(NGSX_RULE_TET_DATA, NGSX_RULE_TET_NPTS);
  CutTable tab;
  for (int dim = 2; dim <= 3; dim++)
    for (int dt = 0; dt < 3; dt++)
      for (int mask = 0; mask < 16; mask++) build_cut_entry(dim, dt, mask, tab.e[dim - 2][dt][mask]);
  DeviceData dd;
  if (cudaMalloc(&dd.rules, all.size() * sizeof(double4)) != cudaSuccess ||
      cudaMalloc(&dd.cut_table, sizeof(CutTable)) != cudaSuccess ||
      cudaMemcpy(dd.rules, all.data(), all.size() * sizeof(double4), cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaMemcpy(dd.cut_table, &tab, sizeof(CutTable), cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaDeviceGetAttribute(&dd.sm_count, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
    set_error(std::string("device_data: ") + cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  // per-call scratch comes from the stream-ordered pool: keep up to 2 GB of it mapped across synchronisation points (the
  // default threshold 0 unmaps everything at every cudaEventSynchronize and re-maps ~100 MB in the next Integrate)
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
    uint64_t keep = 1ull << 31;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  }
  g_dd[dev] = dd;
  g_dd_ready[dev] = true;
  return &g_dd[dev];
}

static thread_local const int32_t *g_packed_output = nullptr;
const int32_t *take_packed_output() {
  const int32_t *p = g_packed_output;
  g_packed_output = nullptr;
  return p;
}

bool rule_lookup(const DeviceData *dd, int sdim, int order, const double4 **ptr, int *n) {
  if (order < 0) order = 0;
  const int *offs = nullptr;
  int maxo = 0;
  switch (sdim) {
    case 0: *ptr = dd->rules; *n = 1; return true;
    case 1: offs = NGSX_RULE_SEGM_OFFS; maxo = NGSX_RULE_SEGM_MAXORDER; break;
    case 2: offs = NGSX_RULE_TRIG_OFFS; maxo = NGSX_RULE_TRIG_MAXORDER; break;
    case 3: offs = NGSX_RULE_TET_OFFS; maxo = NGSX_RULE_TET_MAXORDER; break;
    default: return false;
  }
  if (order > maxo) return false;
  *ptr = dd->rules + g_rule_base[sdim] + offs[order];
  *n = offs[order + 1] - offs[order];
  return true;
}

// sum of the weights of the reference rule of a simplex, in table order (host copy of what the kernels sum)
static double rule_weight_sum(int sdim, int order) {
  if (order < 0) order = 0;
  if (sdim == 0) return 1.0;
  const int *offs = sdim == 1 ? NGSX_RULE_SEGM_OFFS : (sdim == 2 ? NGSX_RULE_TRIG_OFFS : NGSX_RULE_TET_OFFS);
  const double (*data)[4] = sdim == 1 ? NGSX_RULE_SEGM_DATA : (sdim == 2 ? NGSX_RULE_TRIG_DATA : NGSX_RULE_TET_DATA);
  double w = 0.0;
  for (int q = offs[order]; q < offs[order + 1]; q++) w += data[q][3];
  return w;
}

static int check_mesh(const ngsx_mesh *m) {
  NGSX_REQUIRE(m != nullptr, "mesh is NULL");
  NGSX_REQUIRE(m->dim == 2 || m->dim == 3, "mesh.dim must be 2 (triangles) or 3 (tetrahedra)");
  NGSX_REQUIRE(m->ne >= 0 && m->nv >= 0, "negative mesh size");
  NGSX_REQUIRE(m->ne == 0 || (m->vert && m->el2v), "mesh arrays are NULL");
  return NGSX_OK;
}

// =============================================================================================================
// kernels
// =============================================================================================================
template <int DIM>
__device__ __forceinline__ void load_element(const ngsx_mesh &M, const double *__restrict__ lset, int64_t e,
                                             int *v, double *ls, double (*p)[DIM]) {
  if constexpr (DIM == 3) {
    int4 q = __ldg(reinterpret_cast<const int4 *>(M.el2v) + e);
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  } else {
    const int32_t *q = M.el2v + 3 * e;
    v[0] = __ldg(q); v[1] = __ldg(q + 1); v[2] = __ldg(q + 2);
  }
#pragma unroll
  for (int k = 0; k <= DIM; k++) {
    ls[k] = __ldg(lset + v[k]);
    if (p) {
#pragma unroll
      for (int d = 0; d < DIM; d++) p[k][d] = __ldg(M.vert + (int64_t)v[k] * DIM + d);
    }
  }
}

// {count, first element} record of the degenerate-level-set check: reset by a kernel, not by a copy from the host stack, so that
// every entry point of the step can be captured into a CUDA graph
__global__ void k_reset_err(int64_t *err) { err[0] = 0; err[1] = INT64_MAX; }
int reset_err(int64_t *err, cudaStream_t st) {
  k_reset_err<<<1, 1, 0, st>>>(err);
  NGSX_CUDA(cudaGetLastError());
  return NGSX_OK;
}

__global__ void k_clamp(int64_t n, const double *__restrict__ in, double eps, double *__restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double v = in[i];
  if (fabs(v) < eps) v = eps;
  out[i] = v;
}

struct LsetParams { double p[8]; };

__global__ void k_lset_cf(int dim, int64_t nv, const double *__restrict__ vert, int kind, LsetParams P, double eps,
                          double *__restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nv) return;
  double x = vert[i * dim], y = vert[i * dim + 1], z = dim == 3 ? vert[i * dim + 2] : 0.0;
  double v;
  if (kind == NGSX_LSET_AFFINE) v = P.p[0] + P.p[1] * x + P.p[2] * y + P.p[3] * z;
  else {
    double dx = x - P.p[0], dy = y - P.p[1], dz = (dim == 3 && kind != NGSX_LSET_CYLZ) ? z - P.p[2] : 0.0;
    double r = sqrt(dx * dx + dy * dy + dz * dz);
    v = (kind == NGSX_LSET_RING) ? fabs(r - P.p[3]) - P.p[4] : r - P.p[3];
  }
  if (fabs(v) < eps) v = eps;
  out[i] = v;
}

// sum over the sub-simplices of (dt) of the measure factor: the order-0 part volume divided by the rule weight
template <int DIM>
__device__ __forceinline__ double part_factor(const CutEntry &E, const double *ls) {
  double s = 0.0;
  for (int k = 0; k < E.ns; k++) {
    double P[4][3];
    for (int v = 0; v < E.npt; v++) coded_point<DIM>(E.v[k][v], ls, P[v]);
    s += simplex_factor(E.npt, P);
  }
  return s;
}

// (a) classification, two passes:
//   k_classify: persistent CTAs stream the element table (ngsx_stream.cuh); a warp owns four words of every bit array per chunk
//   (ballot, no atomics), issues the 16 level-set gathers of its four elements together, writes domains, the eight bit
//   arrays and the cut ratio of the uncut elements (1 / 0).
//   k_classify_cut: the cut ratio of the IF elements (xfem/cutinfo.cpp:79-91: NEG part / (NEG + POS part); the common rule
//   weight cancels), densely packed from the IF bit array -- fused into the first pass, the sub-simplex walk ran with one or
//   two live lanes in 3 of 4 warps (4 % cut elements: 73 % of the warps hold one) and made the sweep instruction bound.
template <int DIM, bool FULL>
__device__ __forceinline__ void classify_chunk(const ngsx_mesh &M, const double *__restrict__ lset, uint32_t *__restrict__ words,
                                               uint32_t *__restrict__ ifwords, int64_t nwords, int8_t *__restrict__ dom_out,
                                               double *__restrict__ ratio, unsigned long long *__restrict__ err, int64_t c, int64_t first, int n,
                                               const int32_t *stage, int nstaged) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int v[STREAM_WPW][4];
  double ls[STREAM_WPW][4];
#pragma unroll
  for (int k = 0; k < STREAM_WPW; k++) {
    const int i = (warp * STREAM_WPW + k) * 32 + lane;
    if (FULL || i < n) stream_vertices<DIM>(stage, FULL ? STREAM_CH : nstaged, M.el2v, first, i, v[k]);
  }
#pragma unroll
  for (int k = 0; k < STREAM_WPW; k++) {
    const int i = (warp * STREAM_WPW + k) * 32 + lane;
    if (FULL || i < n) {
#pragma unroll
      for (int a = 0; a <= DIM; a++) ls[k][a] = __ldg(lset + v[k][a]);
    }
  }
  unsigned bn[STREAM_WPW], bp[STREAM_WPW], bi[STREAM_WPW], ba[STREAM_WPW];
#pragma unroll
  for (int k = 0; k < STREAM_WPW; k++) {
    const int i = (warp * STREAM_WPW + k) * 32 + lane;
    const int64_t e = first + i;
    int dom = -2;                                    // -2: beyond ne
    if (FULL || i < n) {
      int mask;
      dom = sign_test<DIM>(ls[k], mask);
      if (ratio && dom != IF) ratio[e] = dom == NEG ? 1.0 : (dom == POS ? 0.0 : nan(""));
      if (dom_out) dom_out[e] = (int8_t)dom;
      if (dom == -1 && err) {
        atomicAdd(&err[0], 1ULL);
        atomicMin(reinterpret_cast<long long *>(&err[1]), (long long)e);
      }
    }
    bn[k] = __ballot_sync(0xffffffffu, dom == NEG); bp[k] = __ballot_sync(0xffffffffu, dom == POS);
    bi[k] = __ballot_sync(0xffffffffu, dom == IF); ba[k] = __ballot_sync(0xffffffffu, dom != -2);
  }
  // the warp's 4 words of the 8 bit arrays in one store: lane = array * 4 + word.  The array number is the bit pattern
  // {NEG, POS, IF} of the domains it collects (NGSX_CDOM_*), except ANY = every element
  const int kk = lane & 3, arr = lane >> 2;
  const unsigned n_ = kk == 0 ? bn[0] : kk == 1 ? bn[1] : kk == 2 ? bn[2] : bn[3];
  const unsigned p_ = kk == 0 ? bp[0] : kk == 1 ? bp[1] : kk == 2 ? bp[2] : bp[3];
  const unsigned i_ = kk == 0 ? bi[0] : kk == 1 ? bi[1] : kk == 2 ? bi[2] : bi[3];
  const unsigned a_ = kk == 0 ? ba[0] : kk == 1 ? ba[1] : kk == 2 ? ba[2] : ba[3];
  const unsigned sel = arr == NGSX_CDOM_ANY ? a_ : (((arr & 1) ? n_ : 0u) | ((arr & 2) ? p_ : 0u) | ((arr & 4) ? i_ : 0u));
  const int64_t w = c * (STREAM_CH / 32) + warp * STREAM_WPW + kk;
  if (w < nwords) {
    if (words) words[arr * nwords + w] = sel;
    if (ifwords && arr == NGSX_CDOM_IF) ifwords[w] = sel;
  }
}

template <int DIM>
__global__ void __launch_bounds__(STREAM_THREADS, NGSX_CLS_MINB) k_classify(ngsx_mesh M, const double *__restrict__ lset, uint32_t *__restrict__ words,
                                                               uint32_t *__restrict__ ifwords, int64_t nwords, int8_t *__restrict__ dom_out,
                                                               double *__restrict__ ratio, unsigned long long *__restrict__ err) {
  extern __shared__ __align__(128) unsigned char smem_stream[];
  stream_elements<DIM>(M, smem_stream, [&](int64_t c, int64_t first, int n, const int32_t *stage, int nstaged) {
    if (n == STREAM_CH && nstaged == STREAM_CH) classify_chunk<DIM, true>(M, lset, words, ifwords, nwords, dom_out, ratio, err, c, first, n, stage, nstaged);
    else classify_chunk<DIM, false>(M, lset, words, ifwords, nwords, dom_out, ratio, err, c, first, n, stage, nstaged);
  });
}

constexpr int CUT_G = 4;     // chunks of 32 mask words per CTA of the cut passes

// volume parts only (DIM + 1 points per sub-simplex): unrolled, the points stay in registers
template <int DIM>
__device__ __forceinline__ double part_volume(const CutEntry &E, const double *ls) {
  double s = 0.0;
  const int ns = E.ns;
  for (int k = 0; k < ns; k++) {
    double P[4][3];
#pragma unroll
    for (int v = 0; v <= DIM; v++) coded_point<DIM>(E.v[k][v], ls, P[v]);
    s += simplex_factor(DIM + 1, P);
  }
  return s;
}

// one CTA per CUT_G chunks of 32 IF words (1024 elements each; most chunks hold no cut element); every warp scans the
// chunks, warp w takes the rounds (w - g) mod 8, + 8, ... of 32 items of chunk g; item = (cut element, NEG / POS part) on a lane pair, so
// the two part volumes of an element run side by side.  (Chunks per persistent warp left a 7x imbalance between SMs: the cut
// elements of a sphere cluster where the surface is tangential to the element numbering.)
template <int DIM>
__global__ void __launch_bounds__(256) k_classify_cut(ngsx_mesh M, const double *__restrict__ lset, const CutTable *__restrict__ tab,
                                                     const uint32_t *__restrict__ ifwords, int64_t nwords, double *__restrict__ ratio) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t cwg[CUT_G];
#pragma unroll
  for (int g = 0; g < CUT_G; g++) {
    const int64_t w = ((int64_t)blockIdx.x * CUT_G + g) * 32 + lane;
    cwg[g] = w < nwords ? __ldg(ifwords + w) : 0u;
  }
#pragma unroll
  for (int g = 0; g < CUT_G; g++) {
    const int64_t c = (int64_t)blockIdx.x * CUT_G + g;
    BitDeal D;
    D.scan(cwg[g], lane);
    for (int r0 = 32 * ((warp - g) & 7); r0 < 2 * D.total; r0 += 32 * 8) {
      int owner;
      const int bit = D.item((r0 + lane) >> 1, owner);
      double vol = 0.0;
      int64_t e = 0;
      if (bit >= 0) {
        e = ((c * 32 + owner) << 5) + bit;
        int v[4], mask;
        double ls[4];
        load_element<DIM>(M, lset, e, v, ls, (double(*)[DIM]) nullptr);
        sign_test<DIM>(ls, mask);
        vol = part_volume<DIM>(tab->e[DIM - 2][lane & 1][mask], ls);          // even lane: NEG (0), odd lane: POS (1)
      }
      const double other = __shfl_xor_sync(0xffffffffu, vol, 1);
      if (bit >= 0 && !(lane & 1)) ratio[e] = vol / (vol + other);
    }
  }
}

// boundary elements of a 2D mesh: segments.  Sign test as everywhere; cut ratio = NEG part of the unit segment (the order-0
// rules of xfem/cutinfo.cpp:79-91 on ET_SEGM: the level set is linear, the zero at phi_0 / (phi_0 - phi_1) from vertex 0)
__global__ void __launch_bounds__(256) k_classify_segm(int64_t n, const int32_t *__restrict__ el2v, const double *__restrict__ lset,
                                                      uint32_t *__restrict__ words, int64_t nwords, int8_t *__restrict__ dom_out,
                                                      double *__restrict__ ratio, unsigned long long *__restrict__ err) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int dom = -2;
  if (e < n) {
    const double a = __ldg(lset + __ldg(el2v + 2 * e)), b = __ldg(lset + __ldg(el2v + 2 * e + 1));
    const bool haspos = a > 0.0 || b > 0.0, hasneg = a < 0.0 || b < 0.0, haszero = a == 0.0 || b == 0.0;
    if (haspos && hasneg) dom = haszero ? -1 : IF;
    else if (!haspos && !hasneg) dom = -1;
    else dom = hasneg ? NEG : POS;
    if (dom == -1 && err) {
      atomicAdd(&err[0], 1ULL);
      atomicMin(reinterpret_cast<long long *>(&err[1]), (long long)e);
    }
    if (dom_out) dom_out[e] = (int8_t)dom;
    if (ratio) {
      const double t = a / (a - b);                   // fraction of the segment on vertex 0's side
      ratio[e] = dom == NEG ? 1.0 : (dom == POS ? 0.0 : (dom == IF ? (a < 0.0 ? t : 1.0 - t) : nan("")));
    }
  }
  const unsigned bn = __ballot_sync(0xffffffffu, dom == NEG), bp = __ballot_sync(0xffffffffu, dom == POS);
  const unsigned bi = __ballot_sync(0xffffffffu, dom == IF), ba = __ballot_sync(0xffffffffu, dom != -2);
  const int lane = threadIdx.x & 31;
  const int64_t w = e >> 5;
  if (words && lane < 8 && w < nwords) {
    const unsigned sel = lane == NGSX_CDOM_ANY ? ba : (((lane & 1) ? bn : 0u) | ((lane & 2) ? bp : 0u) | ((lane & 4) ? bi : 0u));
    words[lane * nwords + w] = sel;
  }
}

__global__ void k_node_marks(int64_t ne, int npe, const int32_t *__restrict__ el2node, const int8_t *__restrict__ dom,
                             const double *__restrict__ ratio, int cell_mode, uint32_t *__restrict__ cut_nb, int32_t *__restrict__ dom_of_node) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= ne) return;
  int d = dom[e];
  if (cell_mode) {
    if (d == IF && cut_nb) atomicOr(&cut_nb[e >> 5], 1u << (e & 31));
    if ((d == NEG || d == POS) && dom_of_node) dom_of_node[e] = ratio[e] > 0.5 ? NEG : POS;
    return;
  }
  for (int k = 0; k < npe; k++) {
    int n = el2node[e * npe + k];
    if (d == IF && cut_nb) atomicOr(&cut_nb[n >> 5], 1u << (n & 31));
    if ((d == NEG || d == POS) && dom_of_node) atomicMin(&dom_of_node[n], d);   // integer atomic: deterministic
  }
}

// (b) rule sizes.  uncut: |static rule| if the element lies in dt else 0; cut: ns * |plain rule|
template <int DIM>
__global__ void __launch_bounds__(256) k_rule_count(ngsx_mesh M, const double *__restrict__ lset, const CutTable *__restrict__ tab, int dt,
                                                   int nr_vol, int nr_if, int32_t *__restrict__ npts, unsigned long long *__restrict__ err) {
  __shared__ uint8_t ns_lut[16];
  if (threadIdx.x < 16) ns_lut[threadIdx.x] = tab->e[DIM - 2][dt][threadIdx.x].ns;
  __syncthreads();
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= M.ne) return;
  int v[4], mask;
  double ls[4];
  load_element<DIM>(M, lset, e, v, ls, (double(*)[DIM]) nullptr);
  int dom = sign_test<DIM>(ls, mask);
  int n = 0;
  if (dom == IF) n = ns_lut[mask] * (dt == IF ? nr_if : nr_vol);
  else if (dom == dt) n = nr_vol;
  if (dom == -1 && err) {
    atomicAdd(&err[0], 1ULL);
    atomicMin(reinterpret_cast<long long *>(&err[1]), (long long)e);
  }
  npts[e] = n;
}

// interface weight factor |F^-T n_ref| (TransformQuadUntrafoToIRInterface, straightcutrule.cpp:558-576)
template <int DIM>
__device__ __forceinline__ double interface_scale(const ElGeom<DIM> &G, const double *ls) {
  double g[DIM], n2 = 0.0;
#pragma unroll
  for (int k = 0; k < DIM; k++) { g[k] = ls[k] - ls[DIM]; n2 += g[k] * g[k]; }
  double inv = 1.0 / sqrt(n2), s2 = 0.0;
#pragma unroll
  for (int r = 0; r < DIM; r++) {
    double v = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; k++) v += G.Jinv[k][r] * (g[k] * inv);
    s2 += v * v;
  }
  return sqrt(s2);
}

// (b) emission: one warp per element, lanes stride over the points; order = sub-simplex by sub-simplex, table order inside
template <int DIM>
__global__ void __launch_bounds__(256) k_rule_fill(ngsx_mesh M, const double *__restrict__ lset, const CutTable *__restrict__ tab, int dt,
                                                  const double4 *__restrict__ rule_vol, int nr_vol, const double4 *__restrict__ rule_if, int nr_if,
                                                  int transform_if, int emit_uncut, const int64_t *__restrict__ offs,
                                                  double *__restrict__ pts3, double *__restrict__ wts) {
  __shared__ CutEntry lut[16];
  if (threadIdx.x < 16) lut[threadIdx.x] = tab->e[DIM - 2][dt][threadIdx.x];
  __syncthreads();
  int lane = threadIdx.x & 31;
  int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (e >= M.ne) return;
  int v[4], mask;
  double ls[4], p[DIM + 1][DIM];
  load_element<DIM>(M, lset, e, v, ls, p);
  int dom = sign_test<DIM>(ls, mask);
  int64_t o = offs[e];
  if (dom != IF) {
    if (dom != dt || !emit_uncut) return;
    for (int q = lane; q < nr_vol; q += 32) {
      double4 r = rule_vol[q];
      pts3[3 * (o + q)] = r.x; pts3[3 * (o + q) + 1] = r.y; pts3[3 * (o + q) + 2] = r.z;
      wts[o + q] = r.w;
    }
    return;
  }
  const CutEntry E = lut[mask];
  const double4 *rule = dt == IF ? rule_if : rule_vol;
  int nr = dt == IF ? nr_if : nr_vol;
  double scale = 1.0;
  if (dt == IF && transform_if) {
    ElGeom<DIM> G;
    element_geometry<DIM>(p, G);
    scale = interface_scale<DIM>(G, ls);
  }
  for (int s = 0; s < E.ns; s++) {
    double P[4][3];
    for (int k = 0; k < E.npt; k++) coded_point<DIM>(E.v[s][k], ls, P[k]);
    double fac = simplex_factor(E.npt, P);
    for (int q = lane; q < nr; q += 32) {
      double4 r = rule[q];
      double x[3];
      simplex_point(E.npt, P, r, x);
      int64_t t = o + s * nr + q;
      pts3[3 * t] = x[0]; pts3[3 * t + 1] = x[1]; pts3[3 * t + 2] = x[2];
      wts[t] = r.w * fac * scale;
    }
  }
}

struct PolyDev { int n; double c[NGSX_POLY_MAXTERMS]; int ex[NGSX_POLY_MAXTERMS][3]; };

__device__ __forceinline__ double eval_poly(const PolyDev &f, const double *x) {
  double s = 0.0;
  for (int m = 0; m < f.n; m++) s += f.c[m] * ipow(x[0], f.ex[m][0]) * ipow(x[1], f.ex[m][1]) * ipow(x[2], f.ex[m][2]);
  return s;
}

template <int DIM>
__device__ __forceinline__ void map_point(const double (*p)[DIM], const double *xi, double *x) {
  double lam_last = 1.0;
#pragma unroll
  for (int k = 0; k < DIM; k++) lam_last -= xi[k];
#pragma unroll
  for (int d = 0; d < 3; d++) x[d] = 0.0;
#pragma unroll
  for (int d = 0; d < DIM; d++) {
    double v = lam_last * p[DIM][d];
#pragma unroll
    for (int k = 0; k < DIM; k++) v += xi[k] * p[k][d];
    x[d] = v;
  }
}

// Integrate:
//   1. k_int_uncut / k_int_cut: one thread per element, no block prologue, no atomics -- sign test, then the element's integral:
//      uncut elements of the domain use the static reference rule (first kernel, 10 M threads at full occupancy), cut elements
//      walk their sub-simplices and rule points in a second kernel that only the warps holding a cut element enter (the cut
//      path costs 96 registers: fused, it halved the occupancy of the memory-latency-bound sign pass); constant integrands need
//      no point loop at all (measure factors x sum of the reference weights).  The 32 element integrals of a warp are summed by a butterfly
//      into partial[warp] -- no per-element scratch array, and no block barrier (a warp with cut elements would hold up
//      the seven others of its block: measured 0.32 ms with a block tree against 0.24 ms without).
//   2. k_reduce_final: fixed-shape tree over the warp partials.
// Every partial sum has a shape that depends on ne only: bitwise reproducible, no floating-point atomics.
// phase 1 (lean): sign test, cut mask of every word, integral of the uncut elements of the domain.  Persistent CTAs stream the
// element table (ngsx_stream.cuh); a warp owns four words per chunk and issues the gathers of its four elements together.
template <int DIM, bool FULL>
__device__ __forceinline__ void int_uncut_chunk(const ngsx_mesh &M, const double *__restrict__ lset, int dt, const double4 *__restrict__ rule_vol,
                                                int nr_vol, const PolyDev &f, int fconst, double wref_vol, double *__restrict__ partial,
                                                uint32_t *__restrict__ cutwords, int64_t nred, int64_t c, int64_t first, int n,
                                                const int32_t *stage, int nstaged) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int v[STREAM_WPW][4];
  double ls[STREAM_WPW][4];
#pragma unroll
  for (int k = 0; k < STREAM_WPW; k++) {
    const int i = (warp * STREAM_WPW + k) * 32 + lane;
    if (FULL || i < n) stream_vertices<DIM>(stage, FULL ? STREAM_CH : nstaged, M.el2v, first, i, v[k]);
  }
#pragma unroll
  for (int k = 0; k < STREAM_WPW; k++) {
    const int i = (warp * STREAM_WPW + k) * 32 + lane;
    if (FULL || i < n) {
#pragma unroll
      for (int a = 0; a <= DIM; a++) ls[k][a] = __ldg(lset + v[k][a]);
    }
  }
  double val[STREAM_WPW];
  uint32_t cw[STREAM_WPW];
#pragma unroll
  for (int k = 0; k < STREAM_WPW; k++) {
    const int i = (warp * STREAM_WPW + k) * 32 + lane;
    int dom = -1, mask;
    if (FULL || i < n) dom = sign_test<DIM>(ls[k], mask);
    cw[k] = __ballot_sync(0xffffffffu, dom == IF);
    val[k] = 0.0;
    if (dom == dt && dt != IF) {
      double p[DIM + 1][DIM];
#pragma unroll
      for (int a = 0; a <= DIM; a++)
#pragma unroll
        for (int d = 0; d < DIM; d++) p[a][d] = __ldg(M.vert + (int64_t)v[k][a] * DIM + d);
      ElGeom<DIM> G;
      element_geometry<DIM>(p, G);
      const double meas = fabs(G.det);
      if (fconst) {
        val[k] = meas * (wref_vol * f.c[0]);
      } else {
        double sum = 0.0;
        for (int q = 0; q < nr_vol; q++) {
          const double4 r = rule_vol[q];
          double xi[3] = {r.x, r.y, r.z}, x[3];
          map_point<DIM>(p, xi, x);
          sum += meas * r.w * eval_poly(f, x);
        }
        val[k] = sum;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < STREAM_WPW; k++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) val[k] += __shfl_xor_sync(0xffffffffu, val[k], o);
  }
  if (lane < STREAM_WPW) {
    const int64_t w = c * (STREAM_CH / 32) + warp * STREAM_WPW + lane;
    if (w < nred) {
      partial[w] = lane == 0 ? val[0] : lane == 1 ? val[1] : lane == 2 ? val[2] : val[3];
      cutwords[w] = lane == 0 ? cw[0] : lane == 1 ? cw[1] : lane == 2 ? cw[2] : cw[3];
    }
  }
}

template <int DIM>
__global__ void __launch_bounds__(STREAM_THREADS, NGSX_INT_MINB) k_int_uncut(ngsx_mesh M, const double *__restrict__ lset, int dt, const double4 *__restrict__ rule_vol,
                                                                int nr_vol, PolyDev f, int fconst, double wref_vol, double *__restrict__ partial,
                                                                uint32_t *__restrict__ cutwords, int64_t nred) {
  extern __shared__ __align__(128) unsigned char smem_stream[];
  stream_elements<DIM>(M, smem_stream, [&](int64_t c, int64_t first, int n, const int32_t *stage, int nstaged) {
    if (n == STREAM_CH && nstaged == STREAM_CH) int_uncut_chunk<DIM, true>(M, lset, dt, rule_vol, nr_vol, f, fconst, wref_vol, partial, cutwords, nred, c, first, n, stage, nstaged);
    else int_uncut_chunk<DIM, false>(M, lset, dt, rule_vol, nr_vol, f, fconst, wref_vol, partial, cutwords, nred, c, first, n, stage, nstaged);
  });
  // words beyond the last chunk (the reduction length is padded to a multiple of 8 words)
  const int64_t wend = ((M.ne + STREAM_CH - 1) / STREAM_CH) * (STREAM_CH / 32);
  if (blockIdx.x == 0)
    for (int64_t w = wend + threadIdx.x; w < nred; w += blockDim.x) { partial[w] = 0.0; cutwords[w] = 0u; }
}

// phase 2: the cut elements (1.5 % at config 2), densely packed.  One CTA per CUT_G chunks of 32 mask words (1024 elements
// each; most chunks hold no cut element -- persistent warps with whole chunks each left a 7x imbalance between
// SMs, the cut elements of a sphere cluster where the surface is tangential to the element numbering).  Every warp scans the
// chunks; the set bits are enumerated in (word, bit) order and dealt out 32 at a time, one cut element per lane, warp w taking
// rounds (w - g) mod 8, + 8, ... of chunk g.  The values of one word are summed in bit order by a segmented shuffle loop and handed to the lane that
// owns the word; the per-warp word sums are combined in warp order and added to the word's partial by a single writer.
// Fixed order everywhere: bitwise reproducible, no atomics.
template <int DIM, bool IFACE, bool FCONST>
__global__ void __launch_bounds__(256, FCONST ? 3 : 2) k_int_cut(ngsx_mesh M, const double *__restrict__ lset, const CutTable *__restrict__ tab, int dt,
                                                const double4 *__restrict__ rule_vol, int nr_vol, const double4 *__restrict__ rule_if, int nr_if,
                                                PolyDev f, int fconst, double wref_vol, double wref_if, const uint32_t *__restrict__ cutwords,
                                                int64_t nwords, double *__restrict__ partial) {
  __shared__ double accs[CUT_G][8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t cwg[CUT_G];
  uint32_t mine = 0u;                                // warp g < CUT_G finishes chunk g
#pragma unroll
  for (int g = 0; g < CUT_G; g++) {
    const int64_t w = ((int64_t)blockIdx.x * CUT_G + g) * 32 + lane;
    cwg[g] = w < nwords ? __ldg(cutwords + w) : 0u;
    if (g == warp) mine = cwg[g];
  }
#pragma unroll
  for (int g = 0; g < CUT_G; g++) {
    const int64_t c = (int64_t)blockIdx.x * CUT_G + g;
    BitDeal D;
    D.scan(cwg[g], lane);
    const int total = D.total, begin = D.begin, end = D.end, cnt = D.cnt;
    double acc = 0.0;                                // this warp's share of the cut sum of this lane's word
    for (int r0 = 32 * ((warp - g) & 7); r0 < total; r0 += 32 * 8) {
      int owner;
      const int bit = D.item(r0 + lane, owner);
      double val = 0.0;
      if (bit >= 0) {
        const int64_t e = ((c * 32 + owner) << 5) + bit;
        int v[4], mask;
        double ls[4], p[DIM + 1][DIM];
        load_element<DIM>(M, lset, e, v, ls, p);
        sign_test<DIM>(ls, mask);
        ElGeom<DIM> G;
        element_geometry<DIM>(p, G);
        const double meas = fabs(G.det);
        constexpr int NPT = IFACE ? DIM : DIM + 1;     // points of a sub-simplex: compile time, so they stay in registers
        const CutEntry &E = tab->e[DIM - 2][dt][mask];
        const double4 *rule = IFACE ? rule_if : rule_vol;
        const int nr = IFACE ? nr_if : nr_vol;
        const double scale = IFACE ? interface_scale<DIM>(G, ls) : 1.0;
        const int ns = E.ns;
        for (int sidx = 0; sidx < ns; sidx++) {
          double P[4][3];
#pragma unroll
          for (int k = 0; k < NPT; k++) coded_point<DIM>(E.v[sidx][k], ls, P[k]);
          const double fac = simplex_factor(NPT, P) * scale;
          if constexpr (FCONST) {
            val += meas * (((IFACE ? wref_if : wref_vol) * fac) * f.c[0]);
          } else {
            for (int q = 0; q < nr; q++) {
              double xi[3], x[3];
              const double4 rr = rule[q];
              simplex_point(NPT, P, rr, xi);
              map_point<DIM>(p, xi, x);
              val += meas * (rr.w * fac) * eval_poly(f, x);
            }
          }
        }
      } else {
        owner = 64 + lane;                           // idle lanes: private segment
      }
      // segmented sum in lane order: the head lane of each word's run collects its followers
      double seg = val;
      for (int k = 1; k < 32; k++) {
        const double nv = __shfl_down_sync(0xffffffffu, val, k);
        const int no = __shfl_down_sync(0xffffffffu, owner, k);
        const bool same = (lane + k < 32) && no == owner;
        if (same) seg += nv;
        if (!__any_sync(0xffffffffu, same)) break;
      }
      // word `lane` has items [begin, end) of the chunk; those of this round start at lane max(begin, r0) - r0
      const int firstl = max(begin, r0) - r0;
      const double got = __shfl_sync(0xffffffffu, seg, firstl & 31);
      if (cnt > 0 && begin < r0 + 32 && end > r0) acc += got;
    }
    accs[g][warp][lane] = acc;
  }
  __syncthreads();
  if (warp < CUT_G && mine != 0u) {
    double sum = 0.0;
#pragma unroll
    for (int k = 0; k < 8; k++) sum += accs[warp][k][lane];
    partial[((int64_t)blockIdx.x * CUT_G + warp) * 32 + lane] += sum;
  }
}

template <int DIM, typename... Args>
static void launch_int_cut(bool iface, bool fc, unsigned grid, cudaStream_t st, Args... a) {
  if (iface) { if (fc) k_int_cut<DIM, true, true><<<grid, 256, 0, st>>>(a...); else k_int_cut<DIM, true, false><<<grid, 256, 0, st>>>(a...); }
  else { if (fc) k_int_cut<DIM, false, true><<<grid, 256, 0, st>>>(a...); else k_int_cut<DIM, false, false><<<grid, 256, 0, st>>>(a...); }
}

// ---------------------------------------------------------------------------------------------------------------
// Fused sweep (ngsx_classify_integrate): CutInfo.Update and up to FUSE_MAXINT Integrate calls on the same level set read the
// element table ONCE.  Per chunk the warp does what classify_chunk and int_uncut_chunk do -- sign test, the eight bit arrays,
// domains, cut ratio of the uncut elements, and per integral the static-rule integral of the uncut elements of its domain,
// summed per word by the same butterfly -- so every partial array is bitwise the one ngsx_integrate builds, and the cut passes
// (k_classify_cut, k_int_cut) and the reduction tree run unchanged on top of it.  Three 160 MB streams and three latency
// floors become one (the sweeps are the strong-scaling limiter of the step on a 1/8 mesh).
// ---------------------------------------------------------------------------------------------------------------
constexpr int FUSE_MAXINT = 4;
struct FusedInt {
  int n;
  int dt[FUSE_MAXINT], nr_vol[FUSE_MAXINT], fconst[FUSE_MAXINT];
  const double4 *rule_vol[FUSE_MAXINT];
  double wref_vol[FUSE_MAXINT];
  double *partial[FUSE_MAXINT];
  PolyDev f[FUSE_MAXINT];
};

template <int DIM, bool FULL>
__device__ __forceinline__ void fused_chunk(const ngsx_mesh &M, const double *__restrict__ lset, uint32_t *__restrict__ words,
                                            uint32_t *__restrict__ ifwords, int64_t nwords, int8_t *__restrict__ dom_out,
                                            double *__restrict__ ratio, unsigned long long *__restrict__ err, const FusedInt &I, int64_t nred,
                                            int64_t c, int64_t first, int n, const int32_t *stage, int nstaged) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int v[STREAM_WPW][4];
  double ls[STREAM_WPW][4];
#pragma unroll
  for (int k = 0; k < STREAM_WPW; k++) {
    const int i = (warp * STREAM_WPW + k) * 32 + lane;
    if (FULL || i < n) stream_vertices<DIM>(stage, FULL ? STREAM_CH : nstaged, M.el2v, first, i, v[k]);
  }
#pragma unroll
  for (int k = 0; k < STREAM_WPW; k++) {
    const int i = (warp * STREAM_WPW + k) * 32 + lane;
    if (FULL || i < n) {
#pragma unroll
      for (int a = 0; a <= DIM; a++) ls[k][a] = __ldg(lset + v[k][a]);
    }
  }
  unsigned bn[STREAM_WPW], bp[STREAM_WPW], bi[STREAM_WPW], ba[STREAM_WPW];
  int domk[STREAM_WPW];
#pragma unroll
  for (int k = 0; k < STREAM_WPW; k++) {
    const int i = (warp * STREAM_WPW + k) * 32 + lane;
    const int64_t e = first + i;
    int dom = -2;                                    // -2: beyond ne
    if (FULL || i < n) {
      int mask;
      dom = sign_test<DIM>(ls[k], mask);
      if (ratio && dom != IF) ratio[e] = dom == NEG ? 1.0 : (dom == POS ? 0.0 : nan(""));
      if (dom_out) dom_out[e] = (int8_t)dom;
      if (dom == -1 && err) {
        atomicAdd(&err[0], 1ULL);
        atomicMin(reinterpret_cast<long long *>(&err[1]), (long long)e);
      }
    }
    domk[k] = dom;
    bn[k] = __ballot_sync(0xffffffffu, dom == NEG); bp[k] = __ballot_sync(0xffffffffu, dom == POS);
    bi[k] = __ballot_sync(0xffffffffu, dom == IF); ba[k] = __ballot_sync(0xffffffffu, dom != -2);
  }
  {
    const int kk = lane & 3, arr = lane >> 2;
    const unsigned n_ = kk == 0 ? bn[0] : kk == 1 ? bn[1] : kk == 2 ? bn[2] : bn[3];
    const unsigned p_ = kk == 0 ? bp[0] : kk == 1 ? bp[1] : kk == 2 ? bp[2] : bp[3];
    const unsigned i_ = kk == 0 ? bi[0] : kk == 1 ? bi[1] : kk == 2 ? bi[2] : bi[3];
    const unsigned a_ = kk == 0 ? ba[0] : kk == 1 ? ba[1] : kk == 2 ? ba[2] : ba[3];
    const unsigned sel = arr == NGSX_CDOM_ANY ? a_ : (((arr & 1) ? n_ : 0u) | ((arr & 2) ? p_ : 0u) | ((arr & 4) ? i_ : 0u));
    const int64_t w = c * (STREAM_CH / 32) + warp * STREAM_WPW + kk;
    if (w < nwords) {
      if (words) words[arr * nwords + w] = sel;
      if (ifwords && arr == NGSX_CDOM_IF) ifwords[w] = sel;
    }
  }
  // the uncut parts of the integrals (as int_uncut_chunk)
  double val[FUSE_MAXINT][STREAM_WPW];
#pragma unroll
  for (int j = 0; j < FUSE_MAXINT; j++)
#pragma unroll
    for (int k = 0; k < STREAM_WPW; k++) val[j][k] = 0.0;
#pragma unroll
  for (int k = 0; k < STREAM_WPW; k++) {
    const int dom = domk[k];
    bool need = false;
#pragma unroll
    for (int j = 0; j < FUSE_MAXINT; j++) need |= j < I.n && dom == I.dt[j] && dom != IF;
    if (need) {
      double p[DIM + 1][DIM];
#pragma unroll
      for (int a = 0; a <= DIM; a++)
#pragma unroll
        for (int d = 0; d < DIM; d++) p[a][d] = __ldg(M.vert + (int64_t)v[k][a] * DIM + d);
      ElGeom<DIM> G;
      element_geometry<DIM>(p, G);
      const double meas = fabs(G.det);
#pragma unroll
      for (int j = 0; j < FUSE_MAXINT; j++) {
        if (j < I.n && dom == I.dt[j]) {
          if (I.fconst[j]) {
            val[j][k] = meas * (I.wref_vol[j] * I.f[j].c[0]);
          } else {
            double sum = 0.0;
            for (int q = 0; q < I.nr_vol[j]; q++) {
              const double4 r = I.rule_vol[j][q];
              double xi[3] = {r.x, r.y, r.z}, x[3];
              map_point<DIM>(p, xi, x);
              sum += meas * r.w * eval_poly(I.f[j], x);
            }
            val[j][k] = sum;
          }
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < FUSE_MAXINT; j++) {
    if (j < I.n) {
#pragma unroll
      for (int k = 0; k < STREAM_WPW; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) val[j][k] += __shfl_xor_sync(0xffffffffu, val[j][k], o);
      }
      if (lane < STREAM_WPW) {
        const int64_t w = c * (STREAM_CH / 32) + warp * STREAM_WPW + lane;
        if (w < nred) I.partial[j][w] = lane == 0 ? val[j][0] : lane == 1 ? val[j][1] : lane == 2 ? val[j][2] : val[j][3];
      }
    }
  }
}

template <int DIM>
__global__ void __launch_bounds__(STREAM_THREADS, 2) k_fused_sweep(ngsx_mesh M, const double *__restrict__ lset, uint32_t *__restrict__ words,
                                                                  uint32_t *__restrict__ ifwords, int64_t nwords, int8_t *__restrict__ dom_out,
                                                                  double *__restrict__ ratio, unsigned long long *__restrict__ err,
                                                                  const __grid_constant__ FusedInt I, int64_t nred) {
  extern __shared__ __align__(128) unsigned char smem_stream[];
  stream_elements<DIM>(M, smem_stream, [&](int64_t c, int64_t first, int n, const int32_t *stage, int nstaged) {
    if (n == STREAM_CH && nstaged == STREAM_CH) fused_chunk<DIM, true>(M, lset, words, ifwords, nwords, dom_out, ratio, err, I, nred, c, first, n, stage, nstaged);
    else fused_chunk<DIM, false>(M, lset, words, ifwords, nwords, dom_out, ratio, err, I, nred, c, first, n, stage, nstaged);
  });
  const int64_t wend = ((M.ne + STREAM_CH - 1) / STREAM_CH) * (STREAM_CH / 32);
  if (blockIdx.x == 0)
    for (int j = 0; j < I.n; j++)
      for (int64_t w = wend + threadIdx.x; w < nred; w += blockDim.x) I.partial[j][w] = 0.0;
}

// reduction tree of several integrals at once: blockIdx.y = integral
__global__ void __launch_bounds__(256) k_int_reduce_multi(int64_t n, int64_t stride, const double *__restrict__ in, double *__restrict__ mid, int nmid) {
  __shared__ double wsum[8];
  const double *src = in + (int64_t)blockIdx.y * stride;
  double sum = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) sum += src[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < 8; w++) s += wsum[w];
    mid[(int64_t)blockIdx.y * nmid + blockIdx.x] = s;
  }
}
__global__ void __launch_bounds__(1024) k_reduce_final_multi(int64_t n, const double *__restrict__ mid, double *__restrict__ out) {
  __shared__ double s[1024];
  const double *partial = mid + (int64_t)blockIdx.x * n;
  double acc = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += 1024) acc += partial[i];
  s[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[blockIdx.x] = s[0];
}

// middle level: fixed grid, block-stride over the warp partials, fixed block tree -> one partial per block
__global__ void __launch_bounds__(256) k_int_reduce(int64_t n, const double *__restrict__ in, double *__restrict__ partial) {
  __shared__ double wsum[8];
  double sum = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) sum += in[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < 8; w++) s += wsum[w];
    partial[blockIdx.x] = s;
  }
}

// last pass: one block, fixed association: thread t sums partial[t], partial[t+1024], ... then a fixed tree
__global__ void __launch_bounds__(1024) k_reduce_final(int64_t n, const double *__restrict__ partial, double *__restrict__ out) {
  __shared__ double s[1024];
  double acc = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += 1024) acc += partial[i];
  s[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = s[0];
}

__global__ void k_i32_to_i64(int64_t n, const int32_t *__restrict__ in, int64_t *__restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[i];
}

// facet marking (xfem/cutinfo.cpp:191-255), one thread per facet, ballot -> word
__global__ void __launch_bounds__(256) k_facets_neighbor_types(int64_t nf, const int32_t *__restrict__ f2e, const int32_t *__restrict__ periodic,
                                                              const uint32_t *__restrict__ a,
                                                              const uint32_t *__restrict__ b, int bva, int bvb, int use_and,
                                                              uint32_t *__restrict__ out) {
  int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool r = false;
  if (f < nf) {
    int e0 = f2e[2 * f], e1 = f2e[2 * f + 1];
    bool skip = false;
    if (periodic && e0 >= 0 && e1 < 0) {             // xfem/cutinfo.cpp:225-236: the partner's element, seen from the facet of the smaller number only
      const int f2 = periodic[f];
      if (f2 > f) e1 = f2e[2 * (int64_t)f2];
      else if (f2 >= 0) skip = true;
    }
    if (e0 >= 0 && !skip) {
      bool al = (a[e0 >> 5] >> (e0 & 31)) & 1, bl = (b[e0 >> 5] >> (e0 & 31)) & 1;
      bool ar = e1 >= 0 ? ((a[e1 >> 5] >> (e1 & 31)) & 1) : (bva != 0);
      bool br = e1 >= 0 ? ((b[e1 >> 5] >> (e1 & 31)) & 1) : (bvb != 0);
      r = use_and ? ((al && br) || (ar && bl)) : (al || br || ar || bl);
    }
  }
  unsigned w = __ballot_sync(0xffffffffu, r);
  if ((threadIdx.x & 31) == 0 && (f >> 5) < ((nf + 31) >> 5)) out[f >> 5] = w;
}

__global__ void k_elements_with_neighbor_facets(int64_t nf, const int32_t *__restrict__ f2e, const uint32_t *__restrict__ fw,
                                                uint32_t *__restrict__ out) {
  int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= nf || !((fw[f >> 5] >> (f & 31)) & 1)) return;
  for (int k = 0; k < 2; k++) {
    int e = f2e[2 * f + k];
    if (e >= 0) atomicOr(&out[e >> 5], 1u << (e & 31));
  }
}

__global__ void k_dofs_of_elements(int64_t n, int nd, const int32_t *__restrict__ table, const uint32_t *__restrict__ a,
                                   uint32_t *__restrict__ out) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * nd) return;
  int64_t e = t / nd;
  if (!((a[e >> 5] >> (e & 31)) & 1)) return;
  int d = table[t];
  atomicOr(&out[d >> 5], 1u << (d & 31));
}

}  // namespace ngsx

// =============================================================================================================
// C ABI
// =============================================================================================================
using namespace ngsx;

static inline unsigned nblocks(int64_t n, int bs) { return (unsigned)((n + bs - 1) / bs); }

// ---- ngsx_integrate_static: the allocation-free variant of ngsx_integrate -----------------------------------------------
// Scratch (per-warp partial sums, second reduction level, cut-element words) and the constant tables live in file-scope
// __device__ arrays: every GPU that loads the library holds them (about 1.6 MB) until exit.  In host code these names are
// host-side stand-ins, NOT device addresses: the addresses come from cudaGetSymbolAddress, once per device (static_setup).
constexpr int64_t STATIC_MAX_NE = NGSX_STATIC_MAX_ELEMENTS;
constexpr int64_t STATIC_MAX_NRED = (STATIC_MAX_NE + 255) / 256 * 8;
constexpr int64_t STATIC_NMID = 296;
constexpr size_t STATIC_PARTIAL_BYTES = (size_t)(STATIC_MAX_NRED + STATIC_NMID) * sizeof(double);
constexpr size_t STATIC_SCRATCH_BYTES = STATIC_PARTIAL_BYTES + (size_t)STATIC_MAX_NRED * sizeof(uint32_t);
constexpr int STATIC_NRULES = 1 + NGSX_RULE_SEGM_NPTS + NGSX_RULE_TRIG_NPTS + NGSX_RULE_TET_NPTS;
static __device__ __align__(256) unsigned char s_int_scratch[STATIC_SCRATCH_BYTES];
static __device__ __align__(256) double4 s_int_rules[STATIC_NRULES];
static __device__ __align__(256) CutTable s_int_cut_table;

struct StaticData {
  bool ready = false;
  unsigned char *scratch = nullptr;
  DeviceData dd;                 // rules / cut_table point into the __device__ arrays above
};
static std::mutex g_static_mutex;
static StaticData g_static[64];
static const int s_rule_base[4] = {0, 1, 1 + NGSX_RULE_SEGM_NPTS, 1 + NGSX_RULE_SEGM_NPTS + NGSX_RULE_TRIG_NPTS};

// once per device, with that device current: symbol addresses, constant tables copied in, kernel attributes.  No allocation.
static const StaticData *static_setup() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("cudaGetDevice failed"); return nullptr; }
  std::lock_guard<std::mutex> lock(g_static_mutex);
  StaticData &S = g_static[dev];
  if (S.ready) return &S;
  cudaError_t e = cudaGetSymbolAddress((void **)&S.scratch, s_int_scratch);
  if (e == cudaSuccess) e = cudaGetSymbolAddress((void **)&S.dd.rules, s_int_rules);
  if (e == cudaSuccess) e = cudaGetSymbolAddress((void **)&S.dd.cut_table, s_int_cut_table);
  if (e != cudaSuccess) { set_error(std::string("integrate_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e)); return nullptr; }
  static double4 h_rules[STATIC_NRULES];                           // host staging, fixed size
  static CutTable h_tab;
  int n = 0;
  h_rules[n++] = make_double4(0, 0, 0, 1.0);
  auto push = [&](const double (*d)[4], int cnt) { for (int i = 0; i < cnt; i++) h_rules[n++] = make_double4(d[i][0], d[i][1], d[i][2], d[i][3]); };
  push(NGSX_RULE_SEGM_DATA, NGSX_RULE_SEGM_NPTS);
  push(NGSX_RULE_TRIG_DATA, NGSX_RULE_TRIG_NPTS);
  push(NGSX_RULE_TET_DATA, NGSX_RULE_TET_NPTS);
  for (int dim = 2; dim <= 3; dim++)
    for (int dt = 0; dt < 3; dt++)
      for (int mask = 0; mask < 16; mask++) build_cut_entry(dim, dt, mask, h_tab.e[dim - 2][dt][mask]);
  e = cudaMemcpy(S.dd.rules, h_rules, sizeof(h_rules), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(S.dd.cut_table, &h_tab, sizeof(CutTable), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&S.dd.sm_count, cudaDevAttrMultiProcessorCount, dev);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_int_uncut<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stream_smem<2>());
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_int_uncut<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stream_smem<3>());
  if (e != cudaSuccess) { set_error(std::string("integrate_static set-up: ") + cudaGetErrorString(e)); return nullptr; }
  S.ready = true;
  return &S;
}

static bool static_rule_lookup(const StaticData *S, int sdim, int order, const double4 **ptr, int *n) {
  if (order < 0) order = 0;
  const int *offs = nullptr;
  int maxo = 0;
  switch (sdim) {
    case 0: *ptr = S->dd.rules; *n = 1; return true;
    case 1: offs = NGSX_RULE_SEGM_OFFS; maxo = NGSX_RULE_SEGM_MAXORDER; break;
    case 2: offs = NGSX_RULE_TRIG_OFFS; maxo = NGSX_RULE_TRIG_MAXORDER; break;
    case 3: offs = NGSX_RULE_TET_OFFS; maxo = NGSX_RULE_TET_MAXORDER; break;
    default: return false;
  }
  if (order > maxo) return false;
  *ptr = S->dd.rules + s_rule_base[sdim] + offs[order];
  *n = offs[order + 1] - offs[order];
  return true;
}


// Side streams of the fused sweep: its cut passes (cut ratios, one pass per integral) read the same bit array and write
// disjoint outputs, so they run side by side behind the sweep -- forked from and joined back into the caller's stream with
// events (legal under stream capture: the fork becomes parallel branches of the graph).  Created once per device.
struct ForkPool {
  bool ready = false;
  cudaStream_t s[FUSE_MAXINT];
  cudaEvent_t fork, join[FUSE_MAXINT];
};
static std::mutex g_fork_mutex, g_fork_enqueue_mutex;   // pool creation; one host thread at a time enqueues a fork / join
static ForkPool g_fork[64];
static ForkPool *fork_pool() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(g_fork_mutex);
  ForkPool &F = g_fork[dev];
  if (F.ready) return &F;
  if (cudaEventCreateWithFlags(&F.fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
  for (int i = 0; i < FUSE_MAXINT; i++)
    if (cudaStreamCreateWithFlags(&F.s[i], cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&F.join[i], cudaEventDisableTiming) != cudaSuccess) return nullptr;
  F.ready = true;
  return &F;
}

extern "C" {

void ngsx_set_packed_output(const int32_t *d_rowslot) { ngsx::g_packed_output = d_rowslot; }

int ngsx_abi_version(void) { return NGSX_ABI_VERSION; }
const char *ngsx_last_error(void) { return t_err.c_str(); }
int64_t ngsx_launch_count(void) { return g_launches.load(); }

int ngsx_interpolate_p1_clamp(int64_t nv, const double *in, double eps, double *out, void *stream) {
  NGSX_REQUIRE(nv >= 0 && (nv == 0 || (in && out)), "interpolate_p1_clamp: bad arguments");
  if (nv == 0) return NGSX_OK;
  k_clamp<<<nblocks(nv, 256), 256, 0, as_stream(stream)>>>(nv, in, eps, out);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_interpolate_p1_cf(const ngsx_mesh *mesh, int kind, const double *h_params, double eps, double *out, void *stream) {
  if (int rc = check_mesh(mesh)) return rc;
  NGSX_REQUIRE(kind >= 0 && kind <= NGSX_LSET_CYLZ && h_params && out, "interpolate_p1_cf: bad arguments");
  if (mesh->nv == 0) return NGSX_OK;
  LsetParams P;
  memcpy(P.p, h_params, sizeof P.p);
  k_lset_cf<<<nblocks(mesh->nv, 256), 256, 0, as_stream(stream)>>>(mesh->dim, mesh->nv, mesh->vert, kind, P, eps, out);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_classify(const ngsx_mesh *mesh, const double *lset_v, uint32_t *cdom_words, int8_t *dom, double *ratio,
                  int64_t *err, void *stream) {
  if (int rc = check_mesh(mesh)) return rc;
  NGSX_REQUIRE(lset_v || mesh->ne == 0, "classify: lset_v is NULL");
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  cudaStream_t st = as_stream(stream);
  if (err) { if (int rc = reset_err(err, st)) return rc; }
  if (mesh->ne == 0) return NGSX_OK;
  int64_t nwords = (mesh->ne + 31) / 32;
  auto *e = reinterpret_cast<unsigned long long *>(err);
  NGSX_REQUIRE(mesh->dim == 2 || (reinterpret_cast<uintptr_t>(mesh->el2v) & 15u) == 0, "classify: a 3D mesh.el2v must be 16-byte aligned");
  uint32_t *ifwords = cdom_words ? cdom_words + NGSX_CDOM_IF * nwords : nullptr, *tmp = nullptr;
  if (ratio && !ifwords) {
    NGSX_CUDA(cudaMallocAsync(&tmp, nwords * sizeof(uint32_t), st));
    ifwords = tmp;
  }
  const int64_t nchunks = (mesh->ne + STREAM_CH - 1) / STREAM_CH;
  const unsigned grid = (unsigned)std::min<int64_t>(nchunks, (int64_t)dd->sm_count * NGSX_CLS_MINB);
  const unsigned gcut = (unsigned)(((nwords + 31) / 32 + CUT_G - 1) / CUT_G);      // one CTA per CUT_G chunks of 32 words
  cudaError_t e0 = cudaSuccess;
  if (mesh->dim == 2) {
    static const cudaError_t attr = cudaFuncSetAttribute(k_classify<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stream_smem<2>());
    e0 = attr;
    k_classify<2><<<grid, STREAM_THREADS, stream_smem<2>(), st>>>(*mesh, lset_v, cdom_words, tmp, nwords, dom, ratio, e);
    if (ratio) k_classify_cut<2><<<gcut, 256, 0, st>>>(*mesh, lset_v, dd->cut_table, ifwords, nwords, ratio);
  } else {
    static const cudaError_t attr = cudaFuncSetAttribute(k_classify<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stream_smem<3>());
    e0 = attr;
    k_classify<3><<<grid, STREAM_THREADS, stream_smem<3>(), st>>>(*mesh, lset_v, cdom_words, tmp, nwords, dom, ratio, e);
    if (ratio) k_classify_cut<3><<<gcut, 256, 0, st>>>(*mesh, lset_v, dd->cut_table, ifwords, nwords, ratio);
  }
  cudaError_t e1 = cudaGetLastError();
  g_launches.fetch_add(ratio ? 2 : 1);
  if (tmp) cudaFreeAsync(tmp, st);
  NGSX_CUDA(e0);
  NGSX_CUDA(e1);
  return NGSX_OK;
}

int ngsx_classify_bnd(int dim, int64_t nbe, const int32_t *bnd_el2v, const double *lset_v, uint32_t *cdom_words, int8_t *dom,
                      double *ratio, int64_t *err, void *stream) {
  NGSX_REQUIRE(dim == 2 || dim == 3, "classify_bnd: mesh dimension must be 2 or 3");
  NGSX_REQUIRE(nbe >= 0 && (nbe == 0 || (bnd_el2v && lset_v)), "classify_bnd: bad arguments");
  cudaStream_t st = as_stream(stream);
  if (dim == 3) {                                    // boundary triangles: the sign test and the part volumes need no coordinates
    ngsx_mesh bm{};
    bm.dim = 2; bm.ne = nbe; bm.nv = 0; bm.vert = lset_v; bm.el2v = bnd_el2v;
    return ngsx_classify(&bm, lset_v, cdom_words, dom, ratio, err, stream);
  }
  if (err) { if (int rc = reset_err(err, st)) return rc; }
  if (nbe == 0) return NGSX_OK;
  k_classify_segm<<<nblocks(nbe, 256), 256, 0, st>>>(nbe, bnd_el2v, lset_v, cdom_words, (nbe + 31) / 32, dom, ratio,
                                                      reinterpret_cast<unsigned long long *>(err));
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_node_marks(int64_t ne, int32_t npe, const int32_t *el2node, const int8_t *dom, const double *ratio,
                    int64_t nnodes, int cell_mode, uint32_t *cut_nb_words, int32_t *dom_of_node, void *stream) {
  NGSX_REQUIRE(ne >= 0 && dom && (cell_mode ? ratio != nullptr : (el2node != nullptr && npe > 0)), "node_marks: bad arguments");
  (void)nnodes;
  if (ne == 0) return NGSX_OK;
  k_node_marks<<<nblocks(ne, 256), 256, 0, as_stream(stream)>>>(ne, npe, el2node, dom, ratio, cell_mode, cut_nb_words, dom_of_node);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

static int lookup_rules(const DeviceData *dd, int dim, int order, const double4 **rv, int *nv, const double4 **ri, int *ni) {
  if (!rule_lookup(dd, dim, order, rv, nv) || !rule_lookup(dd, dim - 1, order, ri, ni)) {
    set_error("integration order beyond the tabulated reference rules");
    return NGSX_ERR_UNSUPPORTED;
  }
  return NGSX_OK;
}

int ngsx_rule_count(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, int32_t *npts, int64_t *err, void *stream) {
  if (int rc = check_mesh(mesh)) return rc;
  NGSX_REQUIRE(dt >= 0 && dt <= 2 && npts && lset_v, "rule_count: bad arguments");
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  const double4 *rv, *ri;
  int nv, ni;
  if (int rc = lookup_rules(dd, mesh->dim, order, &rv, &nv, &ri, &ni)) return rc;
  cudaStream_t st = as_stream(stream);
  if (err) { if (int rc = reset_err(err, st)) return rc; }
  if (mesh->ne == 0) return NGSX_OK;
  auto *e = reinterpret_cast<unsigned long long *>(err);
  if (mesh->dim == 2) k_rule_count<2><<<nblocks(mesh->ne, 256), 256, 0, st>>>(*mesh, lset_v, dd->cut_table, dt, nv, ni, npts, e);
  else k_rule_count<3><<<nblocks(mesh->ne, 256), 256, 0, st>>>(*mesh, lset_v, dd->cut_table, dt, nv, ni, npts, e);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_exclusive_scan(int64_t n, const int32_t *in, int64_t *out, void *stream) {
  NGSX_REQUIRE(n >= 0 && out && (n == 0 || in), "exclusive_scan: bad arguments");
  cudaStream_t st = as_stream(stream);
  if (n == 0) { NGSX_CUDA(cudaMemsetAsync(out, 0, sizeof(int64_t), st)); return NGSX_OK; }
  // widen into out[1..n], inclusive-scan in place, out[0] = 0
  k_i32_to_i64<<<nblocks(n, 256), 256, 0, st>>>(n, in, out + 1);
  NGSX_LAUNCH_CHECK();
  NGSX_CUDA(cudaMemsetAsync(out, 0, sizeof(int64_t), st));
  size_t tmp_bytes = 0;
  NGSX_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, out + 1, out + 1, n, st));
  void *tmp = nullptr;
  NGSX_CUDA(cudaMallocAsync(&tmp, tmp_bytes, st));
  cudaError_t e = cub::DeviceScan::InclusiveSum(tmp, tmp_bytes, out + 1, out + 1, n, st);
  g_launches.fetch_add(2);
  cudaFreeAsync(tmp, st);
  NGSX_CUDA(e);
  return NGSX_OK;
}

int ngsx_rule_fill(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, int transform_if, int emit_uncut,
                   const int64_t *offs, double *pts3, double *wts, void *stream) {
  if (int rc = check_mesh(mesh)) return rc;
  NGSX_REQUIRE(dt >= 0 && dt <= 2 && offs && pts3 && wts && lset_v, "rule_fill: bad arguments");
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  const double4 *rv, *ri;
  int nv, ni;
  if (int rc = lookup_rules(dd, mesh->dim, order, &rv, &nv, &ri, &ni)) return rc;
  if (mesh->ne == 0) return NGSX_OK;
  cudaStream_t st = as_stream(stream);
  unsigned nb = nblocks(mesh->ne * 32, 256);
  if (mesh->dim == 2) k_rule_fill<2><<<nb, 256, 0, st>>>(*mesh, lset_v, dd->cut_table, dt, rv, nv, ri, ni, transform_if, emit_uncut, offs, pts3, wts);
  else k_rule_fill<3><<<nb, 256, 0, st>>>(*mesh, lset_v, dd->cut_table, dt, rv, nv, ri, ni, transform_if, emit_uncut, offs, pts3, wts);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_classify_integrate(const ngsx_mesh *mesh, const double *lset_v, uint32_t *cdom_words, int8_t *dom, double *ratio, int64_t *err,
                            int32_t nint, const int32_t *h_dt, const int32_t *h_order, const ngsx_poly *h_f, double *out, void *stream) {
  if (int rc = check_mesh(mesh)) return rc;
  NGSX_REQUIRE(lset_v || mesh->ne == 0, "classify_integrate: lset_v is NULL");
  NGSX_REQUIRE(nint >= 0 && nint <= FUSE_MAXINT && (nint == 0 || (h_dt && h_order && h_f && out)), "classify_integrate: bad integrals (at most 4)");
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  cudaStream_t st = as_stream(stream);
  if (err) { if (int rc = reset_err(err, st)) return rc; }
  if (mesh->ne == 0) { if (nint) NGSX_CUDA(cudaMemsetAsync(out, 0, nint * sizeof(double), st)); return NGSX_OK; }
  NGSX_REQUIRE(mesh->dim == 2 || (reinterpret_cast<uintptr_t>(mesh->el2v) & 15u) == 0, "classify_integrate: a 3D mesh.el2v must be 16-byte aligned");
  const int64_t ne = mesh->ne, nwords = (ne + 31) / 32;
  const int64_t nred = (int64_t)nblocks(ne, 256) * 8;                 // the reduction shape of ngsx_integrate
  const unsigned nmid = (unsigned)std::min<int64_t>((nred + 255) / 256, 296);
  FusedInt I{};
  I.n = nint;
  const double4 *ri[FUSE_MAXINT] = {nullptr, nullptr, nullptr, nullptr};
  int ni[FUSE_MAXINT] = {0, 0, 0, 0};
  double wref_if[FUSE_MAXINT] = {0.0, 0.0, 0.0, 0.0};
  for (int j = 0; j < nint; j++) {
    NGSX_REQUIRE(h_dt[j] >= 0 && h_dt[j] <= 2 && h_f[j].nterms >= 0 && h_f[j].nterms <= NGSX_POLY_MAXTERMS, "classify_integrate: bad integral");
    if (int rc = lookup_rules(dd, mesh->dim, h_order[j], &I.rule_vol[j], &I.nr_vol[j], &ri[j], &ni[j])) return rc;
    I.dt[j] = h_dt[j];
    I.f[j].n = h_f[j].nterms;
    for (int m = 0; m < I.f[j].n; m++) { I.f[j].c[m] = h_f[j].coef[m]; for (int d = 0; d < 3; d++) I.f[j].ex[m][d] = h_f[j].ex[m][d]; }
    I.fconst[j] = I.f[j].n == 1 && I.f[j].ex[0][0] == 0 && I.f[j].ex[0][1] == 0 && I.f[j].ex[0][2] == 0;
    I.wref_vol[j] = rule_weight_sum(mesh->dim, h_order[j]);
    wref_if[j] = rule_weight_sum(mesh->dim - 1, h_order[j]);
  }
  double *partial = nullptr;
  uint32_t *tmp = nullptr;
  if (nint) NGSX_CUDA(cudaMallocAsync(&partial, (size_t)nint * (nred + nmid) * sizeof(double), st));
  uint32_t *ifwords = cdom_words ? cdom_words + NGSX_CDOM_IF * nwords : nullptr;
  if (!ifwords) { NGSX_CUDA(cudaMallocAsync(&tmp, nwords * sizeof(uint32_t), st)); ifwords = tmp; }
  double *mid = partial ? partial + (size_t)nint * nred : nullptr;
  for (int j = 0; j < nint; j++) I.partial[j] = partial + (size_t)j * nred;
  const int64_t nchunks = (ne + STREAM_CH - 1) / STREAM_CH;
  const unsigned nb = (unsigned)std::min<int64_t>(nchunks, (int64_t)dd->sm_count * 2);
  const unsigned gcut = (unsigned)(((nwords + 31) / 32 + CUT_G - 1) / CUT_G);
  auto *e = reinterpret_cast<unsigned long long *>(err);
  cudaError_t e0 = cudaSuccess;
  int launches = 1;
  // the cut passes: the first on the caller's stream, the others on side streams forked behind the sweep
  const int ntasks = (ratio ? 1 : 0) + nint;
  ForkPool *F = ntasks > 1 ? fork_pool() : nullptr;
  std::unique_lock<std::mutex> enqueue(g_fork_enqueue_mutex, std::defer_lock);
  if (F) enqueue.lock();
  int task = 0;
  cudaError_t ef = cudaSuccess;
  auto after_sweep = [&]() { if (F) ef = cudaEventRecord(F->fork, st); };
  auto next_stream = [&]() -> cudaStream_t {
    const int t = task++;
    if (!F || t == 0) return st;
    if (ef == cudaSuccess) ef = cudaStreamWaitEvent(F->s[t - 1], F->fork, 0);
    return F->s[t - 1];
  };
  if (mesh->dim == 2) {
    static const cudaError_t attr = cudaFuncSetAttribute(k_fused_sweep<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stream_smem<2>());
    e0 = attr;
    k_fused_sweep<2><<<nb, STREAM_THREADS, stream_smem<2>(), st>>>(*mesh, lset_v, cdom_words, tmp, nwords, dom, ratio, e, I, nred);
    after_sweep();
    if (ratio) { k_classify_cut<2><<<gcut, 256, 0, next_stream()>>>(*mesh, lset_v, dd->cut_table, ifwords, nwords, ratio); launches++; }
    for (int j = 0; j < nint; j++)
      launch_int_cut<2>(I.dt[j] == IF, I.fconst[j], gcut, next_stream(), *mesh, lset_v, dd->cut_table, I.dt[j], I.rule_vol[j], I.nr_vol[j], ri[j], ni[j],
                        I.f[j], I.fconst[j], I.wref_vol[j], wref_if[j], (const uint32_t *)ifwords, nwords, I.partial[j]);
  } else {
    static const cudaError_t attr = cudaFuncSetAttribute(k_fused_sweep<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stream_smem<3>());
    e0 = attr;
    k_fused_sweep<3><<<nb, STREAM_THREADS, stream_smem<3>(), st>>>(*mesh, lset_v, cdom_words, tmp, nwords, dom, ratio, e, I, nred);
    after_sweep();
    if (ratio) { k_classify_cut<3><<<gcut, 256, 0, next_stream()>>>(*mesh, lset_v, dd->cut_table, ifwords, nwords, ratio); launches++; }
    for (int j = 0; j < nint; j++)
      launch_int_cut<3>(I.dt[j] == IF, I.fconst[j], gcut, next_stream(), *mesh, lset_v, dd->cut_table, I.dt[j], I.rule_vol[j], I.nr_vol[j], ri[j], ni[j],
                        I.f[j], I.fconst[j], I.wref_vol[j], wref_if[j], (const uint32_t *)ifwords, nwords, I.partial[j]);
  }
  cudaError_t e1 = cudaGetLastError();
  for (int t = 1; F && t < task; t++) {                                // join
    if (ef == cudaSuccess) ef = cudaEventRecord(F->join[t - 1], F->s[t - 1]);
    if (ef == cudaSuccess) ef = cudaStreamWaitEvent(st, F->join[t - 1], 0);
  }
  if (nint) {
    k_int_reduce_multi<<<dim3(nmid, nint), 256, 0, st>>>(nred, nred, partial, mid, (int)nmid);
    k_reduce_final_multi<<<nint, 1024, 0, st>>>(nmid, mid, out);
    launches += nint + 2;
  }
  g_launches.fetch_add(launches);
  cudaError_t e2 = cudaGetLastError();
  if (partial) cudaFreeAsync(partial, st);
  if (tmp) cudaFreeAsync(tmp, st);
  NGSX_CUDA(e0);
  NGSX_CUDA(e1);
  NGSX_CUDA(e2);
  NGSX_CUDA(ef);
  return NGSX_OK;
}

int ngsx_integrate(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, const ngsx_poly *h_f, double *out, void *stream) {
  if (int rc = check_mesh(mesh)) return rc;
  NGSX_REQUIRE(dt >= 0 && dt <= 2 && h_f && out && lset_v, "integrate: bad arguments");
  NGSX_REQUIRE(h_f->nterms >= 0 && h_f->nterms <= NGSX_POLY_MAXTERMS, "integrate: too many polynomial terms");
  const DeviceData *dd = device_data();
  if (!dd) return NGSX_ERR_CUDA;
  const double4 *rv, *ri;
  int nv, ni;
  if (int rc = lookup_rules(dd, mesh->dim, order, &rv, &nv, &ri, &ni)) return rc;
  cudaStream_t st = as_stream(stream);
  if (mesh->ne == 0) { NGSX_CUDA(cudaMemsetAsync(out, 0, sizeof(double), st)); return NGSX_OK; }
  PolyDev f;
  f.n = h_f->nterms;
  for (int m = 0; m < f.n; m++) { f.c[m] = h_f->coef[m]; for (int d = 0; d < 3; d++) f.ex[m][d] = h_f->ex[m][d]; }
  const int fconst = f.n == 1 && f.ex[0][0] == 0 && f.ex[0][1] == 0 && f.ex[0][2] == 0;
  const double wref_vol = rule_weight_sum(mesh->dim, order), wref_if = rule_weight_sum(mesh->dim - 1, order);
  const int64_t ne = mesh->ne;
  const int64_t nred = (int64_t)nblocks(ne, 256) * 8;       // one partial per warp of 32 elements: device-independent reduction shape
  double *partial = nullptr;
  const unsigned nmid = (unsigned)std::min<int64_t>((nred + 255) / 256, 296);   // fixed upper bound: device-independent reduction shape
  uint32_t *cutwords = nullptr;
  NGSX_CUDA(cudaMallocAsync(&partial, (nred + nmid) * sizeof(double), st));
  NGSX_CUDA(cudaMallocAsync(&cutwords, nred * sizeof(uint32_t), st));
  NGSX_REQUIRE(mesh->dim == 2 || (reinterpret_cast<uintptr_t>(mesh->el2v) & 15u) == 0, "integrate: a 3D mesh.el2v must be 16-byte aligned");
  const int64_t nchunks = (ne + STREAM_CH - 1) / STREAM_CH;
  const unsigned nb = (unsigned)std::min<int64_t>(nchunks, (int64_t)dd->sm_count * NGSX_INT_MINB);   // persistent
  const unsigned ncut = (unsigned)(((nred + 31) / 32 + CUT_G - 1) / CUT_G);                    // one CTA per CUT_G chunks of 32 words
  cudaError_t e0;
  if (mesh->dim == 2) {
    static const cudaError_t attr = cudaFuncSetAttribute(k_int_uncut<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stream_smem<2>());
    e0 = attr;
    k_int_uncut<2><<<nb, STREAM_THREADS, stream_smem<2>(), st>>>(*mesh, lset_v, dt, rv, nv, f, fconst, wref_vol, partial, cutwords, nred);
    launch_int_cut<2>(dt == IF, fconst, ncut, st, *mesh, lset_v, dd->cut_table, dt, rv, nv, ri, ni, f, fconst, wref_vol, wref_if, cutwords, nred, partial);
  } else {
    static const cudaError_t attr = cudaFuncSetAttribute(k_int_uncut<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stream_smem<3>());
    e0 = attr;
    k_int_uncut<3><<<nb, STREAM_THREADS, stream_smem<3>(), st>>>(*mesh, lset_v, dt, rv, nv, f, fconst, wref_vol, partial, cutwords, nred);
    launch_int_cut<3>(dt == IF, fconst, ncut, st, *mesh, lset_v, dd->cut_table, dt, rv, nv, ri, ni, f, fconst, wref_vol, wref_if, cutwords, nred, partial);
  }
  cudaError_t e1 = cudaGetLastError();
  k_int_reduce<<<nmid, 256, 0, st>>>(nred, partial, partial + nred);
  k_reduce_final<<<1, 1024, 0, st>>>(nmid, partial + nred, out);
  g_launches.fetch_add(4);
  cudaError_t e2 = cudaGetLastError();
  cudaFreeAsync(partial, st); cudaFreeAsync(cutwords, st);
  NGSX_CUDA(e0);
  NGSX_CUDA(e1);
  NGSX_CUDA(e2);
  return NGSX_OK;
}

int ngsx_facets_neighbor_types(int64_t nf, const int32_t *facet2el, const uint32_t *a_words, const uint32_t *b_words,
                               int bnd_val_a, int bnd_val_b, int use_and, uint32_t *out_words, void *stream) {
  NGSX_REQUIRE(nf >= 0 && (nf == 0 || (facet2el && a_words && b_words && out_words)), "facets_neighbor_types: bad arguments");
  if (nf == 0) return NGSX_OK;
  k_facets_neighbor_types<<<nblocks(nf, 256), 256, 0, as_stream(stream)>>>(nf, facet2el, nullptr, a_words, b_words, bnd_val_a, bnd_val_b, use_and, out_words);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_facets_neighbor_types_periodic(int64_t nf, const int32_t *facet2el, const int32_t *periodic_facet, const uint32_t *a_words,
                                        const uint32_t *b_words, int bnd_val_a, int bnd_val_b, int use_and, uint32_t *out_words, void *stream) {
  NGSX_REQUIRE(nf >= 0 && (nf == 0 || (facet2el && periodic_facet && a_words && b_words && out_words)), "facets_neighbor_types_periodic: bad arguments");
  if (nf == 0) return NGSX_OK;
  k_facets_neighbor_types<<<nblocks(nf, 256), 256, 0, as_stream(stream)>>>(nf, facet2el, periodic_facet, a_words, b_words, bnd_val_a, bnd_val_b, use_and, out_words);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_elements_with_neighbor_facets(int64_t nf, const int32_t *facet2el, const uint32_t *facet_words,
                                       int64_t ne, uint32_t *out_words, void *stream) {
  NGSX_REQUIRE(nf >= 0 && ne >= 0 && (nf == 0 || (facet2el && facet_words)) && (ne == 0 || out_words), "elements_with_neighbor_facets: bad arguments");
  cudaStream_t st = as_stream(stream);
  if (ne > 0) NGSX_CUDA(cudaMemsetAsync(out_words, 0, ((ne + 31) / 32) * sizeof(uint32_t), st));
  if (nf == 0) return NGSX_OK;
  k_elements_with_neighbor_facets<<<nblocks(nf, 256), 256, 0, st>>>(nf, facet2el, facet_words, out_words);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

int ngsx_dofs_of_elements(int64_t n, int32_t nd, const int32_t *table, const uint32_t *a_words, int64_t ndof, uint32_t *out_words, void *stream) {
  NGSX_REQUIRE(n >= 0 && nd > 0 && ndof >= 0 && (n == 0 || (table && a_words)) && (ndof == 0 || out_words), "dofs_of_elements: bad arguments");
  cudaStream_t st = as_stream(stream);
  if (ndof > 0) NGSX_CUDA(cudaMemsetAsync(out_words, 0, ((ndof + 31) / 32) * sizeof(uint32_t), st));
  if (n == 0) return NGSX_OK;
  k_dofs_of_elements<<<nblocks(n * nd, 256), 256, 0, st>>>(n, nd, table, a_words, out_words);
  NGSX_LAUNCH_CHECK();
  return NGSX_OK;
}

// Allocation-free ngsx_integrate: same kernels, same launch shapes, same reduction tree (bitwise the same result); the scratch
// is the static device array.  All calls on a device share it: calls that overlap on the GPU (different streams) race.  It
// keeps the last call's contents; nothing has to start at zero -- k_int_uncut writes every partial and every cut word it or
// the later kernels read.  No cudaMalloc / cudaFree / pool allocation on this path.
int ngsx_integrate_static(const ngsx_mesh *mesh, const double *lset_v, int dt, int order, const ngsx_poly *h_f, double *out, void *stream) {
  if (int rc = check_mesh(mesh)) return rc;
  NGSX_REQUIRE(dt >= 0 && dt <= 2 && h_f && out && lset_v, "integrate_static: bad arguments");
  NGSX_REQUIRE(h_f->nterms >= 0 && h_f->nterms <= NGSX_POLY_MAXTERMS, "integrate_static: too many polynomial terms");
  if (mesh->ne > STATIC_MAX_NE) {
    set_error("integrate_static: mesh.ne = " + std::to_string((long long)mesh->ne) + " exceeds NGSX_STATIC_MAX_ELEMENTS = " +
              std::to_string((long long)STATIC_MAX_NE) + " (the static scratch array); use ngsx_integrate");
    return NGSX_ERR_CAPACITY;
  }
  const StaticData *S = static_setup();
  if (!S) return NGSX_ERR_CUDA;
  const double4 *rv, *ri;
  int nv, ni;
  if (!static_rule_lookup(S, mesh->dim, order, &rv, &nv) || !static_rule_lookup(S, mesh->dim - 1, order, &ri, &ni)) {
    set_error("integration order beyond the tabulated reference rules");
    return NGSX_ERR_UNSUPPORTED;
  }
  cudaStream_t st = as_stream(stream);
  if (mesh->ne == 0) { NGSX_CUDA(cudaMemsetAsync(out, 0, sizeof(double), st)); return NGSX_OK; }
  NGSX_REQUIRE(mesh->dim == 2 || (reinterpret_cast<uintptr_t>(mesh->el2v) & 15u) == 0, "integrate_static: a 3D mesh.el2v must be 16-byte aligned");
  PolyDev f;
  f.n = h_f->nterms;
  for (int m = 0; m < f.n; m++) { f.c[m] = h_f->coef[m]; for (int d = 0; d < 3; d++) f.ex[m][d] = h_f->ex[m][d]; }
  const int fconst = f.n == 1 && f.ex[0][0] == 0 && f.ex[0][1] == 0 && f.ex[0][2] == 0;
  const double wref_vol = rule_weight_sum(mesh->dim, order), wref_if = rule_weight_sum(mesh->dim - 1, order);
  const int64_t ne = mesh->ne;
  const int64_t nred = (int64_t)nblocks(ne, 256) * 8;
  const unsigned nmid = (unsigned)std::min<int64_t>((nred + 255) / 256, STATIC_NMID);
  double *partial = reinterpret_cast<double *>(S->scratch);                                  // [nred] + [nmid]
  uint32_t *cutwords = reinterpret_cast<uint32_t *>(S->scratch + STATIC_PARTIAL_BYTES);      // [nred]
  const int64_t nchunks = (ne + STREAM_CH - 1) / STREAM_CH;
  const unsigned nb = (unsigned)std::min<int64_t>(nchunks, (int64_t)S->dd.sm_count * NGSX_INT_MINB);
  const unsigned ncut = (unsigned)(((nred + 31) / 32 + CUT_G - 1) / CUT_G);
  if (mesh->dim == 2) {
    k_int_uncut<2><<<nb, STREAM_THREADS, stream_smem<2>(), st>>>(*mesh, lset_v, dt, rv, nv, f, fconst, wref_vol, partial, cutwords, nred);
    launch_int_cut<2>(dt == IF, fconst, ncut, st, *mesh, lset_v, S->dd.cut_table, dt, rv, nv, ri, ni, f, fconst, wref_vol, wref_if, cutwords, nred, partial);
  } else {
    k_int_uncut<3><<<nb, STREAM_THREADS, stream_smem<3>(), st>>>(*mesh, lset_v, dt, rv, nv, f, fconst, wref_vol, partial, cutwords, nred);
    launch_int_cut<3>(dt == IF, fconst, ncut, st, *mesh, lset_v, S->dd.cut_table, dt, rv, nv, ri, ni, f, fconst, wref_vol, wref_if, cutwords, nred, partial);
  }
  cudaError_t e1 = cudaGetLastError();
  k_int_reduce<<<nmid, 256, 0, st>>>(nred, partial, partial + nred);
  k_reduce_final<<<1, 1024, 0, st>>>(nmid, partial + nred, out);
  g_launches.fetch_add(4);
  cudaError_t e2 = cudaGetLastError();
  NGSX_CUDA(e1);
  NGSX_CUDA(e2);
  return NGSX_OK;
}

}  // extern "C"
