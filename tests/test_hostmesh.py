"""The numpy set-up of the CPU arm (oracle/hostmesh.py) numbers vertices, elements, edges and dofs exactly like the
device-side builders (ngsxfem_b200/meshes.py), so the two arms of bench.py describe the same problem and the full-size
parity gate can compare CSR values entry by entry.  Also: the C helpers for the sparsity pattern and the colouring."""
import numpy as np
import pytest
import torch

from ngsxfem_b200 import meshes
from oracle import hostmesh, orc


@pytest.mark.parametrize("dim,nx", [(2, 7), (3, 5)])
def test_same_numbering_as_device_builders(dim, nx):
    m = meshes.MakeStructured2DMesh(nx) if dim == 2 else meshes.MakeStructured3DMesh(nx)
    vert, el2v = hostmesh.structured_2d(nx) if dim == 2 else hostmesh.structured_3d(nx)
    assert np.array_equal(m.el2v.numpy(), el2v) and np.array_equal(m.vert.numpy(), vert)
    for order in (1, 2):
        e2d, ndof = m.h1_dofs(order)
        he2d, hndof = hostmesh.h1_dofs(el2v, vert.shape[0], order)
        assert ndof == hndof and np.array_equal(e2d.numpy(), he2d)


def test_pattern_and_colouring():
    import scipy.sparse as sp
    vert, el2v = hostmesh.structured_3d(4)
    e2d, ndof = hostmesh.h1_dofs(el2v, vert.shape[0], 2)
    els = np.arange(0, el2v.shape[0], 3)
    tab = e2d[els]
    rowptr, colidx = hostmesh.csr_pattern(tab, ndof)
    nd = tab.shape[1]
    rows = np.repeat(tab, nd, axis=1).reshape(-1); cols = np.tile(tab, (1, nd)).reshape(-1)
    patt = sp.csr_matrix((np.ones(len(rows)), (rows, cols)), shape=(ndof, ndof))
    patt.sum_duplicates(); patt.sort_indices()
    assert np.array_equal(patt.indptr, rowptr) and np.array_equal(patt.indices, colidx)
    perm, cptr = hostmesh.color_elements(tab, ndof)
    assert sorted(perm.tolist()) == list(range(len(els))) and cptr[0] == 0 and cptr[-1] == len(els)
    for c in range(len(cptr) - 1):                           # no two elements of one colour share a dof
        d = tab[perm[cptr[c]:cptr[c + 1]]].reshape(-1)
        assert len(np.unique(d)) == len(d)


def test_reference_problem_small():
    """the CPU arm's step on a small mesh: compiled reference (or the port) against the plain oracle"""
    import bench
    R = bench.ReferenceProblem(6)
    secs, res = R.step(1)
    dom, ratio, nerr = orc.cutinfo_mesh(3, R.el2v, R.vert, R.ls)
    assert nerr == 0 and np.array_equal(dom, res["dom"])
    cptr = np.array([0, len(R.active)], dtype=np.int64)
    vals, nass = orc.assemble_csr(3, R.el2v, R.vert, R.ls, 2, R.e2d, 0, 4, 1.0, 1.0, cptr, R.active.astype(np.int32), R.rowptr, R.colidx, nthreads=1)
    assert nass == res["nass"]
    par = bench.parity_record(dict(dom=dom, ratio=ratio, vals=vals, s_if=orc.integrate_mesh(3, R.el2v, R.vert, R.ls, 2, 4)[0],
                                   s_neg=orc.integrate_mesh(3, R.el2v, R.vert, R.ls, 0, 4)[0]), res, R.rowptr)
    assert par["ok"], par
    s2, r2 = R.step(3)                                       # the bounded sample keeps the colouring valid
    assert r2["n_classified"] == (R.ne + 2) // 3 and 0 < r2["nass"] <= res["nass"]
