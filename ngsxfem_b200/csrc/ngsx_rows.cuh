// ngsx_rows.cuh -- the owner-computes ("row gather") assembly plan shared by ngsx_rows.cu (plan construction, row kernel),
// ngsx_rowsmat.cu (gather of precomputed blocks) and ngsx_assemble.cu (geometry / cut-element passes that feed them).
//
// Plan = a static schedule.  The non-empty CSR rows of the plan are cut into TILES of consecutive rows whose values fit a warp's
// shared-memory window; inside a tile every row belongs to one LANE (longest-processing-time assignment: rows by decreasing pair
// count to the least loaded lane, lanes then numbered by decreasing load), and the lane walks the (row, element) pairs of its
// rows one per STEP.  The 16-byte pair descriptors are stored step-major, lane-minor and without padding: the lanes active in
// step s are the prefix [0, cnt(s)), so the warp reads one contiguous run of descriptors per step (coalesced 128-bit loads).
// A descriptor carries everything the lane needs -- no per-row records, no split tables:
//   word 0      k | i << 27      list index of the element (27 bits) and local row i (5 bits)
//   word 1      off | pos[0] << 16 | pos[1] << 24      offset of the row inside the tile's window; positions of the local columns
//   words 2, 3  pos[2..9]        (32-byte descriptors, 10 < nd <= 26: four more words with pos[10..25])
// The same numbering addresses PACKED block rows (ngsx_rowplan_rowslots): slot = index of the pair's descriptor.
#pragma once
#include "ngsx_host.cuh"

namespace ngsx {
struct TileHdr {             // 64 bytes
  int64_t z0;                // first CSR entry of the tile's span
  int64_t d0;                // first descriptor (= slot) of the tile
  int32_t zlen, nsteps;      // CSR entries in the span; steps = pairs of the most loaded lane
  uint8_t load[32];          // pairs per lane, non-increasing
  int32_t pad[2];
};
static_assert(sizeof(TileHdr) == 64, "tile header layout");
constexpr uint32_t DESC_KBITS = 27;
constexpr uint32_t DESC_KMASK = (1u << DESC_KBITS) - 1u;

// moments of a permutation-symmetric reference rule over the barycentrics (host-computed, kernel argument):
// m0 = sum w, m1 = sum w l_a, m2s = sum w l_a^2, m2d = sum w l_a l_b, m3aab, m3abc, m4aabb, m4aabc, m4abcd
struct RuleMom { double m0, m1, m2s, m2d, m3aab, m3abc, m4aabb, m4aabc, m4abcd; };
}  // namespace ngsx

struct ngsx_rowplan {
  int64_t nlist = 0, ndof = 0, npairs = 0, nrows = 0, nnz_covered = 0, max_row = 0, nnz_lo = 0, nnz_hi = 0;
  int32_t nd = 0, ds = 0;
  int32_t lower = 0;                // plan over a lower-triangle (symmetric storage) pattern: columns above the row carry POS_SKIP
  int64_t ntiles = 0;               // non-empty tiles
  int64_t nslots = 0;               // == npairs (no padding)
  int64_t max_steps = 0;
  ngsx::TileHdr *tiles = nullptr;   // [ntiles] in processing order
  uint8_t *desc = nullptr;          // [nslots][ds]
  int32_t *rowslot = nullptr;       // [nlist][nd] slot of pair (k, i) or -1; built on first request
  const int64_t *rowptr = nullptr;  // caller-owned CSR (must outlive the plan)
  // scratch of the numeric phase of ngsx_assemble_csr_rows_h1, allocated on first use
  double *G = nullptr;              // [nlist][dim+1][4] per local vertex a: K[a][0..dim-1], (pad), |det J| c_mass  -- or a NaN-boxed code
  int32_t *cutlist = nullptr;       // [nlist] list indices of the elements that need quadrature
  int32_t *cutcount = nullptr;      // [1] work-list length of the last call (may exceed cut_cap: the overflow is reported)
  unsigned *counter = nullptr;      // [1] tile claim counter of the row kernels
  double *cutmats = nullptr;        // [cut_cap][nd*nd]
  int64_t cut_cap = 0;
  int g_dim = 0;
};

namespace ngsx {

// tuning knobs (overridable with -D for sweeps)
#ifndef NGSX_RW
#define NGSX_RW 4
#endif
#ifndef NGSX_ACCW
#define NGSX_ACCW 1400
#endif
#ifndef NGSX_TILE_MAXROWS
#define NGSX_TILE_MAXROWS 64
#endif
#ifndef NGSX_MINBLOCKS
#define NGSX_MINBLOCKS 4
#endif
#ifndef NGSX_ROWS_UNROLL
#define NGSX_ROWS_UNROLL 2
#endif
#ifndef NGSX_ROWS_COSTORDER_BYTES
#define NGSX_ROWS_COSTORDER_BYTES (64ll << 20)   // plans whose geometry records fit here: tiles ordered longest first
#endif
#ifndef NGSX_TILE_MAXPAIRS
#define NGSX_TILE_MAXPAIRS 768
#endif
constexpr int RW = NGSX_RW;                        // warps per CTA of the row kernels
constexpr int MAXROW = 256;                        // longest supported CSR row (column positions are bytes)
constexpr int POS_SKIP = 255;                      // lower-triangle plans: this local column has no entry in the row (col > row); rows < 255 entries
constexpr int ACCW = NGSX_ACCW;                    // doubles of shared memory per warp: the CSR span of one tile
constexpr int TILE_MAXROWS = NGSX_TILE_MAXROWS;
constexpr int TILE_MAXPAIRS = NGSX_TILE_MAXPAIRS;

// NaN-boxed code in the mass slot of a geometry record: quiet NaN | payload; payload 0 = the element contributes nothing,
// s + 1 = row s of cutmats
constexpr unsigned long long GCODE_NAN = 0x7ff8000000000000ULL;

// software prefetch (no destination register): the schedule is static, so a warp can pull the descriptor block of its tile into
// L2 when it starts the tile, and the geometry record of a pair into L1 two steps before it is used
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// predicated 128-bit load INTO d: the registers keep their value when p is false.  (An `if (p) d = __ldg(..)` may compile to a
// load into a temporary and a predicated move right behind it, whose scoreboard wait exposes the full latency of the load.)
__device__ __forceinline__ void ldg128_if(bool p, const void *ptr, uint4 &d) {
  asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %4, 0;\n\t@q ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%5];\n\t}"
               : "+r"(d.x), "+r"(d.y), "+r"(d.z), "+r"(d.w) : "r"((uint32_t)p), "l"(ptr));
}
__device__ __forceinline__ void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
// the tile's contiguous block of `count` records of `bytes` bytes starting at record d0, 128-byte lines dealt to the lanes
__device__ __forceinline__ void prefetch_block_l2(const void *base, int64_t d0, int count, int bytes, int lane) {
  const char *p = reinterpret_cast<const char *>(base) + d0 * bytes;
  const int total = count * bytes;
  for (int off = lane * 128; off < total; off += 32 * 128) prefetch_l2(p + off);
}

int rows_scratch(ngsx_rowplan *plan, int dim);
int rows_reserve_cut(ngsx_rowplan *plan, int64_t cap);
// moments of the reference rule of a dim-simplex; false if the rule is not permutation symmetric (or not tabulated)
bool rule_moments(int dim, int order, RuleMom *out);
int launch_rows(const ngsx_rowplan *plan, int dim, const RuleMom &mom, int accumulate, double *vals, int sm_count, cudaStream_t st);

}  // namespace ngsx
