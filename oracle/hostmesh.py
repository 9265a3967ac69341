"""TEST INFRASTRUCTURE -- numpy set-up of the benchmark problem for the CPU arm (bench.py --impl reference, the
cpu_baseline leg and the full-size parity gate).  No torch, no CUDA, no product library: the reference arm has to run
on a GPU-less host.  Not part of the product path.

Same numbering as ngsolve.meshes.MakeStructured2DMesh / MakeStructured3DMesh (NGS/python/meshes.py:75-121, 259-313) and as
ngsxfem_b200/meshes.py (edges = sorted vertex pairs in lexicographic order), so element e / dof d here are element e / dof d
of the device-side problem and the two CSR value arrays can be compared entry by entry."""
import ctypes

import numpy as np

from . import orc

TRIG_EDGES = ((2, 0), (1, 2), (0, 1))                                   # NGS/fem/elementtopology.cpp:143-146
TET_EDGES = ((3, 0), (3, 1), (3, 2), (0, 1), (0, 2), (1, 2))            # NGS/fem/elementtopology.cpp:161-167
KUHN_TETS = ((0, 4, 5, 6), (0, 6, 7, 4), (0, 3, 7, 6), (0, 1, 6, 5), (0, 1, 2, 6), (0, 3, 6, 2))   # NGS/python/meshes.py:307-313


def structured_3d(nx, ny=None, nz=None):
    """(vert [nv][3] f64, el2v [ne][4] i32): Kuhn tets, vertex id = i (ny+1)(nz+1) + j (nz+1) + k, cube-major elements"""
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    i, j, k = np.meshgrid(np.arange(nx + 1) / nx, np.arange(ny + 1) / ny, np.arange(nz + 1) / nz, indexing="ij")
    vert = np.stack([i.ravel(), j.ravel(), k.ravel()], axis=1)
    ci, cj, ck = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    base = (ci * (ny + 1) * (nz + 1) + cj * (nz + 1) + ck).ravel().astype(np.int64)
    up = base + (ny + 1) * (nz + 1)
    pnum = np.stack([base, base + 1, base + (nz + 1) + 1, base + (nz + 1), up, up + 1, up + (nz + 1) + 1, up + (nz + 1)], axis=1)
    el2v = pnum[:, np.array(KUHN_TETS)].reshape(-1, 4).astype(np.int32)
    return np.ascontiguousarray(vert), np.ascontiguousarray(el2v)


def structured_2d(nx, ny=None):
    ny = nx if ny is None else ny
    yy, xx = np.meshgrid(np.arange(ny + 1) / ny, np.arange(nx + 1) / nx, indexing="ij")
    vert = np.stack([xx.ravel(), yy.ravel()], axis=1)
    ci, cj = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    base = (ci * (nx + 1) + cj).ravel().astype(np.int64)
    t1 = np.stack([base, base + 1, base + nx + 1], axis=1)
    t2 = np.stack([base + 1, base + nx + 2, base + nx + 1], axis=1)
    el2v = np.stack([t1, t2], axis=1).reshape(-1, 3).astype(np.int32)
    return np.ascontiguousarray(vert), np.ascontiguousarray(el2v)


def edges(el2v, nv):
    """(edge2v [nedge][2], el2edge [ne][nel]): edges numbered by their sorted vertex pair, lexicographically"""
    loc = TRIG_EDGES if el2v.shape[1] == 3 else TET_EDGES
    ev = el2v.astype(np.int64)
    a = np.stack([ev[:, p] for p, _ in loc], axis=1)
    b = np.stack([ev[:, q] for _, q in loc], axis=1)
    key = (np.minimum(a, b) * nv + np.maximum(a, b)).ravel()
    ukey, inv = np.unique(key, return_inverse=True)
    return np.stack([ukey // nv, ukey % nv], axis=1), inv.reshape(el2v.shape[0], len(loc))


def h1_dofs(el2v, nv, order):
    """H1HighOrderFESpace numbering for order 1..2 (NGS/comp/h1hofespace.cpp:562-572, 1123-1154): vertices, then edges"""
    assert order in (1, 2)
    if order == 1:
        return np.ascontiguousarray(el2v.astype(np.int32)), nv
    e2v, el2e = edges(el2v, nv)
    return np.ascontiguousarray(np.concatenate([el2v.astype(np.int64), nv + el2e], axis=1).astype(np.int32)), nv + e2v.shape[0]


def sphere_p1(vert, center, radius, eps=1e-14):
    """InterpolateToP1 of |x - c| - r with the clamp of utils/p1interpol.cpp:88-90.  Same operation order as the device
    kernel (k_lset_cf): sum of squares in x, y, z order, sqrt, subtract -- every step correctly rounded on both sides."""
    d = vert - np.asarray(center, dtype=np.float64)
    s = d[:, 0] * d[:, 0]
    for c in range(1, vert.shape[1]):
        s = s + d[:, c] * d[:, c]
    v = np.sqrt(s) - radius
    v[np.abs(v) < eps] = eps
    return v


def csr_pattern(table, ndof):
    table = np.ascontiguousarray(table, dtype=np.int32)
    n, nd = table.shape
    L = orc.load()
    rowptr = np.zeros(ndof + 1, dtype=np.int64)
    nnz = L.orc_csr_pattern(n, nd, orc.P(table, orc.ip), ndof, orc.P(rowptr, orc.lp), None)
    colidx = np.zeros(nnz, dtype=np.int32)
    L.orc_csr_pattern(n, nd, orc.P(table, orc.ip), ndof, orc.P(rowptr, orc.lp), orc.P(colidx, orc.ip))
    return rowptr, colidx


def color_elements(table, ndof):
    """(perm, color_ptr): greedy element colouring (NGS/comp/fespace.cpp:598-705 in spirit: no two elements of a colour share a dof)"""
    table = np.ascontiguousarray(table, dtype=np.int32)
    n, nd = table.shape
    perm = np.zeros(n, dtype=np.int64)
    cptr = np.zeros(130, dtype=np.int64)
    nc = orc.load().orc_color_elements(n, nd, orc.P(table, orc.ip), ndof, orc.P(perm, orc.lp), orc.P(cptr, orc.lp))
    if nc < 0:
        raise RuntimeError("more than 128 colours")
    return perm, cptr[: nc + 1].copy()


def config2(nx, radius=0.35):
    """BASELINE.json configs[1]: Kuhn mesh of the unit cube, sphere level set, P2 H1, HASNEG restriction"""
    vert, el2v = structured_3d(nx)
    ls = sphere_p1(vert, (0.5, 0.5, 0.5), radius)
    e2d, ndof = h1_dofs(el2v, vert.shape[0], 2)
    return dict(dim=3, vert=vert, el2v=el2v, ls=ls, el2dof=e2d, ndof=ndof)
