// ngsx_rows.cuh -- the owner-computes ("row gather") assembly plan shared by ngsx_rows.cu (plan construction, row kernel)
// and ngsx_assemble.cu (geometry-factor / cut-element passes that feed it).
#pragma once
#include "ngsx_host.cuh"

// One (row, element) pair descriptor, 16 bytes (one LDG.128):
//   int32 k        list index of the element
//   uint8 i        local row (the dof of the row inside element k)
//   uint8 pos[nd]  position of local column j inside the CSR row (nd <= 11; 32-byte descriptors for 11 < nd <= 27:
//                  the merged dof lists of facet patches)
namespace ngsx {
struct RowRec { uint32_t p0; uint16_t np; uint16_t off; };   // first pair, number of pairs, offset of the row inside its tile
struct TileInfo { int64_t z0; int32_t zlen; int32_t pad; };  // CSR span of a tile
}
struct ngsx_rowplan {
  int64_t nlist = 0, ndof = 0, npairs = 0, nrows = 0, nnz_covered = 0, max_row = 0, nnz_lo = 0, nnz_hi = 0;
  int32_t nd = 0, ds = 0;
  int32_t *rows = nullptr;          // [nrows] dof number of the non-empty rows, ascending
  int64_t *pairptr = nullptr;       // [nrows + 1]
  uint8_t *desc = nullptr;          // [npairs][ds]
  ngsx::RowRec *rowrec = nullptr;   // [nrows]
  int32_t *tile_q = nullptr;        // [ntiles + 1] first plan row of each tile (natural order)
  ngsx::TileInfo *tile_info = nullptr;  // [ntiles_active] in processing order
  int32_t *tile_split = nullptr;    // [ntiles_active][33] first plan row of each lane's row range (pair-balanced)
  int64_t ntiles = 0, ntiles_active = 0;
  const int64_t *rowptr = nullptr;  // caller-owned CSR (must outlive the plan)
  // scratch of the numeric phase, allocated on first use
  double *G = nullptr;              // [nlist][8] geometry factors + code
  int32_t *cutlist = nullptr;       // [nlist] list indices of the elements that need quadrature
  int32_t *cutcount = nullptr;      // [1]
  double *cutmats = nullptr;        // [nlist][nd*nd]
  double *tab = nullptr;            // [(npair+1)][nd*nd] reference tables of the constant-rule path
  const void *tab_rule = nullptr;   // rule the tables were computed for (they depend on the reference rule only: computed once)
  int tab_key = 0;
};

namespace ngsx {

// tuning knobs (overridable with -D for sweeps; the defaults are the measured best on B200 at config 2)
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
#ifndef NGSX_TILE_MAXPAIRS
#define NGSX_TILE_MAXPAIRS 768
#endif
constexpr int RW = NGSX_RW;                        // warps per CTA of the row kernel
constexpr int MAXROW = 256;                        // longest supported CSR row (column positions are bytes)
constexpr int ACCW = NGSX_ACCW;                    // doubles of shared memory per warp: the CSR span of one tile (11 KB -> 4 CTAs = 16 warps per SM)
constexpr int TILE_MAXROWS = NGSX_TILE_MAXROWS;    // rows per tile (2 per lane) ...
constexpr int TILE_MAXPAIRS = NGSX_TILE_MAXPAIRS;  // ... and (row, element) pairs per tile (24 per lane)

__device__ __forceinline__ RowRec ld_rowrec(const RowRec *p) {
  const uint2 u = __ldg(reinterpret_cast<const uint2 *>(p));
  RowRec r;
  r.p0 = u.x; r.np = (uint16_t)(u.y & 0xffffu); r.off = (uint16_t)(u.y >> 16);
  return r;
}

// reference tables of the constant-rule path (k_ref_tables -> k_rows):  [0, nd^2) mass table Tm[i][j];  then per shape i the
// gradient-moment block cM[i][a][n] (dim x (dim+1) doubles) at stride CS -- CS/2 is odd so that the 16-byte shared-memory
// slots of different i fall into different bank groups
template <int DIM> struct RefTab {
  static constexpr int CS = DIM == 3 ? 14 : 6;
  static __host__ __device__ constexpr int cm_offset(int ne) { return (ne + 1) & ~1; }    // 16-byte aligned start of the cM blocks
};

// geometry-factor record (8 doubles per list entry): g[0..5] = |det J| c_stiff (J^-1 J^-T) upper triangle, g[6] = |det J| c_mass,
// g[7] = code as int64: -1 no contribution, 0 constant-rule element, s+1 = row s of cutmats
constexpr int GSTRIDE = 8;

int rows_scratch(ngsx_rowplan *plan, int npairtab);
int launch_rows(const ngsx_rowplan *plan, int npairtab, int accumulate, double *vals, int sm_count, cudaStream_t st);

}  // namespace ngsx
