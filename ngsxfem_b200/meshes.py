"""Synthetic simplex meshes + H1 dof tables (host-side plumbing; torch is used as the array library so the
10M-tet tables can be built on the GPU).

Mirrors ``ngsolve.meshes.MakeStructured2DMesh(quads=False)`` / ``MakeStructured3DMesh(hexes=False)``
(NGS/python/meshes.py:75-121 and :259-313): same vertex numbering, same element order, same local vertex
order -- so element number ``e`` here is element ``e`` of the reference mesh.

Dof numbering follows ``H1HighOrderFESpace`` (NGS/comp/h1hofespace.cpp:562-572, 1123-1154): vertex dofs, then
``order-1`` dofs per edge, then face dofs, then cell dofs; element-local order = vertices, local edges
(NGS/fem/elementtopology.cpp:143-146 trig, :161-167 tet), local faces (:171-176), cell.  The global *edge/face
enumeration* is ours (sorted vertex tuples); parity is per element matrix, so any fixed enumeration is valid
(SURVEY.md section 8d).
"""
from dataclasses import dataclass, field
from typing import Optional

import torch

TRIG_EDGES = ((2, 0), (1, 2), (0, 1))
TET_EDGES = ((3, 0), (3, 1), (3, 2), (0, 1), (0, 2), (1, 2))
TET_FACES = ((3, 1, 2), (3, 2, 0), (3, 0, 1), (0, 2, 1))
KUHN_TETS = ((0, 4, 5, 6), (0, 6, 7, 4), (0, 3, 7, 6), (0, 1, 6, 5), (0, 1, 2, 6), (0, 3, 6, 2))


@dataclass
class SimplexMesh:
    """Flat-array simplex mesh: ``vert[nv][dim]`` f64, ``el2v[ne][dim+1]`` i32 (the C-ABI layout)."""
    dim: int
    vert: torch.Tensor
    el2v: torch.Tensor
    _cache: dict = field(default_factory=dict, repr=False)

    @property
    def ne(self):
        return self.el2v.shape[0]

    @property
    def nv(self):
        return self.vert.shape[0]

    @property
    def device(self):
        return self.vert.device

    def to(self, device):
        return SimplexMesh(self.dim, self.vert.to(device), self.el2v.to(device))

    # ---- topology tables ---------------------------------------------------------------------------
    def _unique_rows(self, rows):
        """rows: [n, k] int64 sorted within row -> (unique rows, inverse)"""
        nv = self.nv
        key = rows[:, 0]
        for c in range(1, rows.shape[1]):
            key = key * nv + rows[:, c]
        ukey, inv = torch.unique(key, return_inverse=True)
        cols = []
        k = ukey
        for c in range(rows.shape[1] - 1, 0, -1):
            cols.append(k % nv)
            k = k // nv
        cols.append(k)
        return torch.stack(cols[::-1], dim=1), inv

    def edges(self):
        """(edge2v [nedge,2] i64 sorted, el2edge [ne, nedge_local] i64)"""
        if "edges" not in self._cache:
            loc = TRIG_EDGES if self.dim == 2 else TET_EDGES
            ev = self.el2v.long()
            pairs = torch.stack([torch.stack([ev[:, a], ev[:, b]], dim=1) for a, b in loc], dim=1)   # ne, nel, 2
            pairs = torch.sort(pairs, dim=2).values.reshape(-1, 2)
            uniq, inv = self._unique_rows(pairs)
            self._cache["edges"] = (uniq, inv.reshape(self.ne, len(loc)))
        return self._cache["edges"]

    def faces(self):
        """3D: (face2v [nface,3] sorted, el2face [ne,4]).  2D: faces are the elements."""
        if self.dim == 2:
            return None, torch.arange(self.ne, device=self.device).reshape(-1, 1)
        if "faces" not in self._cache:
            ev = self.el2v.long()
            tri = torch.stack([torch.stack([ev[:, a], ev[:, b], ev[:, c]], dim=1) for a, b, c in TET_FACES], dim=1)
            tri = torch.sort(tri, dim=2).values.reshape(-1, 3)
            uniq, inv = self._unique_rows(tri)
            self._cache["faces"] = (uniq, inv.reshape(self.ne, 4))
        return self._cache["faces"]

    def facets(self):
        """(el2facet [ne, nfl] i64, facet2el [nfacet,2] i32 with -1 for 'no neighbour'; lower element number first)"""
        if "facets" not in self._cache:
            if self.dim == 2:
                _, el2f = self.edges()
            else:
                _, el2f = self.faces()
            nf = int(el2f.max().item()) + 1
            nfl = el2f.shape[1]
            elid = torch.arange(self.ne, device=self.device).repeat_interleave(nfl)
            flat = el2f.reshape(-1)
            order = torch.sort(flat * self.ne + elid).values       # facet-major, element ascending
            f_sorted = order // self.ne
            e_sorted = order % self.ne
            first = torch.ones_like(f_sorted, dtype=torch.bool)
            first[1:] = f_sorted[1:] != f_sorted[:-1]
            f2e = torch.full((nf, 2), -1, dtype=torch.int32, device=self.device)
            f2e[f_sorted[first], 0] = e_sorted[first].int()
            f2e[f_sorted[~first], 1] = e_sorted[~first].int()
            self._cache["facets"] = (el2f, f2e)
        return self._cache["facets"]

    def boundary_elements(self):
        """el2v [nbe][dim] int32 of the boundary elements (facets with one neighbour; NGSolve's surface / segment elements),
        in facet order"""
        if "bnd" not in self._cache:
            _, f2e = self.facets()
            f2v = self.edges()[0] if self.dim == 2 else self.faces()[0]
            self._cache["bnd"] = f2v[f2e[:, 1] < 0].to(torch.int32).contiguous()
        return self._cache["bnd"]

    # ---- H1 dof table ------------------------------------------------------------------------------
    def h1_dofs(self, order):
        """(el2dof [ne, ndof_el] i32, ndof) for H1 of the given order (1..3)."""
        key = ("h1", order)
        if key in self._cache:
            return self._cache[key]
        assert 1 <= order <= 3
        cols = [self.el2v.long()]
        ndof = self.nv
        if order >= 2:
            e2v, el2e = self.edges()
            ne_dofs = order - 1
            base = ndof + el2e * ne_dofs
            cols.append(torch.stack([base + k for k in range(ne_dofs)], dim=2).reshape(self.ne, -1))
            ndof += e2v.shape[0] * ne_dofs
        if order >= 3:
            if self.dim == 2:
                cols.append(ndof + torch.arange(self.ne, device=self.device).reshape(-1, 1))
                ndof += self.ne
            else:
                f2v, el2f = self.faces()
                cols.append(ndof + el2f)
                ndof += f2v.shape[0]
        el2dof = torch.cat(cols, dim=1).to(torch.int32).contiguous()
        self._cache[key] = (el2dof, ndof)
        return self._cache[key]


@dataclass
class TensorMesh:
    """Flat-array mesh of quadrilaterals / hexahedra: ``el2v[ne][2^dim]`` in NGSolve's local vertex order ((0,0), (1,0), (1,1),
    (0,1) [, the same at the upper face]); SURVEY section 8 row f-5.  Only what the cut-quadrature path needs (P1 level sets)."""
    dim: int
    vert: torch.Tensor
    el2v: torch.Tensor
    tensor = True

    @property
    def ne(self):
        return self.el2v.shape[0]

    @property
    def nv(self):
        return self.vert.shape[0]

    @property
    def device(self):
        return self.vert.device

    def h1_dofs(self, order):
        assert order == 1, "tensor-product cells carry the (multilinear) level set only"
        return self.el2v, self.nv


def MakeStructured2DMesh(nx=10, ny=None, device="cpu", lo=(0.0, 0.0), hi=(1.0, 1.0), quads=False):
    """Triangles of NGS/python/meshes.py:75-121 (quads=False): vertex id = i*(nx+1)+j at (j/nx, i/ny);
    per cell trigs [base, base+1, base+nx+1] and [base+1, base+nx+2, base+nx+1].  ``lo/hi`` apply the affine
    ``mapping`` the reference allows (e.g. the [-1,1]^2 square of fictdom_ghostpen.py)."""
    ny = nx if ny is None else ny
    i = torch.arange(ny + 1, device=device, dtype=torch.float64)
    j = torch.arange(nx + 1, device=device, dtype=torch.float64)
    yy, xx = torch.meshgrid(i / ny, j / nx, indexing="ij")
    vert = torch.stack([xx.reshape(-1), yy.reshape(-1)], dim=1)
    if tuple(lo) != (0.0, 0.0) or tuple(hi) != (1.0, 1.0):
        lo_t = torch.tensor(lo, device=device, dtype=torch.float64)
        hi_t = torch.tensor(hi, device=device, dtype=torch.float64)
        vert = lo_t + vert * (hi_t - lo_t)
    ci, cj = torch.meshgrid(torch.arange(ny, device=device), torch.arange(nx, device=device), indexing="ij")
    base = (ci * (nx + 1) + cj).reshape(-1)
    if quads:                                       # NGS/python/meshes.py:113-119: pnum = [base, base+1, base+nx+2, base+nx+1]
        el2v = torch.stack([base, base + 1, base + nx + 2, base + nx + 1], dim=1).to(torch.int32).contiguous()
        return TensorMesh(2, vert.contiguous(), el2v)
    t1 = torch.stack([base, base + 1, base + nx + 1], dim=1)
    t2 = torch.stack([base + 1, base + nx + 2, base + nx + 1], dim=1)
    el2v = torch.stack([t1, t2], dim=1).reshape(-1, 3).to(torch.int32).contiguous()
    return SimplexMesh(2, vert.contiguous(), el2v)


def MakeStructured3DMesh(nx=10, ny=None, nz=None, device="cpu", lo=(0.0, 0.0, 0.0), hi=(1.0, 1.0, 1.0), hexes=False):
    """Kuhn tets of NGS/python/meshes.py:259-313 (hexes=False): vertex id = i*(ny+1)*(nz+1)+j*(nz+1)+k at
    (i/nx, j/ny, k/nz); 6 tets per cube in the reference's order; element order is cube-major."""
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    i = torch.arange(nx + 1, device=device, dtype=torch.float64) / nx
    j = torch.arange(ny + 1, device=device, dtype=torch.float64) / ny
    k = torch.arange(nz + 1, device=device, dtype=torch.float64) / nz
    X, Y, Z = torch.meshgrid(i, j, k, indexing="ij")
    vert = torch.stack([X.reshape(-1), Y.reshape(-1), Z.reshape(-1)], dim=1)
    if tuple(lo) != (0.0, 0.0, 0.0) or tuple(hi) != (1.0, 1.0, 1.0):
        lo_t = torch.tensor(lo, device=device, dtype=torch.float64)
        hi_t = torch.tensor(hi, device=device, dtype=torch.float64)
        vert = lo_t + vert * (hi_t - lo_t)
    ci, cj, ck = torch.meshgrid(torch.arange(nx, device=device), torch.arange(ny, device=device),
                                torch.arange(nz, device=device), indexing="ij")
    base = (ci * (ny + 1) * (nz + 1) + cj * (nz + 1) + ck).reshape(-1)
    up = base + (ny + 1) * (nz + 1)
    pnum = torch.stack([base, base + 1, base + (nz + 1) + 1, base + (nz + 1),
                        up, up + 1, up + (nz + 1) + 1, up + (nz + 1)], dim=1)
    if hexes:                                       # NGS/python/meshes.py:296-304: the cube's eight corners in pnum order
        return TensorMesh(3, vert.contiguous(), pnum.to(torch.int32).contiguous())
    q = torch.tensor(KUHN_TETS, device=device, dtype=torch.long)          # 6 x 4
    el2v = pnum[:, q].reshape(-1, 4).to(torch.int32).contiguous()
    return SimplexMesh(3, vert.contiguous(), el2v)


def sphere_levelset(vert: torch.Tensor, center, radius: float) -> torch.Tensor:
    """phi = |x - c| - r evaluated at the vertices (the InterpolateToP1 CF branch, utils/p1interpol.cpp:37-72,
    evaluates the CF at the vertex position)."""
    c = torch.tensor(center, device=vert.device, dtype=torch.float64)
    return torch.sqrt(((vert - c) ** 2).sum(dim=1)) - radius
