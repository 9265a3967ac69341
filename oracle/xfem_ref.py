"""TEST INFRASTRUCTURE -- CPU restatement of the reference's XFEM element matrices (numpy, small cases only).

Follows the reference directly, not the CUDA decomposition:
  * XFESpace::Update (xfem/xFESpace.cpp:244-366): x-dofs = base dofs of IF elements in base-dof order, domofdof = NEG unless
    the node of the dof touches an uncut element, then INVERT(domain of that element) (dom_of_node: xfem/cutinfo.cpp:152-186);
  * DiffOpX<D, RNEG/RPOS/RNEG_GRAD/RPOS_GRAD>::GenerateMatrix (xfem/xfemdiffops.cpp:8-66): B-row i = base shape (gradient) if
    xsign[i] == domain of the operator, else 0;
  * the two-phase Nitsche forms of py_tutorials/nxfem.py:77-107 evaluated point by point on the cut rules
    (xfem/symboliccutbfi.cpp:73-315; intorder = 2 order - difforders, :97-107).
The cut rules and H1 shapes come from the pinned C oracle (oracle/ngsx_oracle.c, checked against the compiled reference and
the reference's known-answer tests).  Parity of the mask algebra itself is pinned only by this restatement: "parity unpinned"
for the XFEM row in the sense of the task statement (no golden vector of the reference exists for it; the reference's
Python bindings do not build here)."""
import numpy as np

from . import orc

NEG, POS, IF = 0, 1, 2


def xfespace(dim, order, el2v, el2dof, ndof_base, nv, dom):
    """returns (basedof2xdof [ndof_base], domofdof [nx]) -- xfem/xFESpace.cpp:276-366 for an H1 space with vertex + edge dofs"""
    active = np.zeros(ndof_base, dtype=bool)
    for e in np.nonzero(dom == IF)[0]:
        active[el2dof[e]] = True
    b2x = np.where(active, np.cumsum(active) - 1, -1)
    # dom_of_node per base dof (vertex dofs: vertex nodes; edge dofs: edge nodes): IF unless adjacent to an uncut element
    node_dom = np.full(ndof_base, IF, dtype=np.int64)
    for e in np.nonzero(dom != IF)[0]:
        node_dom[el2dof[e]] = dom[e]
    domofdof = np.full(int(active.sum()), NEG, dtype=np.int64)
    for b in np.nonzero(active)[0]:
        if node_dom[b] != IF:
            domofdof[b2x[b]] = 1 - node_dom[b]
    return b2x, domofdof


def _bvectors(dim, order, vnums, pmat, xi, xsign, nrm):
    """B-vectors of the proxies at one point over [dofs of Vh | x-dofs]: values U[k], gradients G[k] (k = NEG, POS)"""
    s, ds = orc.h1_shape(dim, order, vnums, xi[:dim])
    J = (pmat[:dim] - pmat[dim]).T
    g = ds @ np.linalg.inv(J)
    nd = len(s)
    U, G = [], []
    for k in (NEG, POS):
        m = (np.asarray(xsign) == k).astype(float)
        U.append(np.concatenate([s, m * s]))
        G.append(np.vstack([g, m[:, None] * g]))
    return U, G


def nxfem_elmat(dim, order, vnums, pmat, ls, xsign, kappa_neg, alpha, stab_factor):
    """compound element matrix (2 nd x 2 nd) of the three integrators of py_tutorials/nxfem.py:98-107 on one element:
    NEG: alpha0 gradu0.gradv0, POS: alpha1 gradu1.gradv1, IF: avgflux_u [v] + avgflux_v [u] + stab [u][v],
    stab = stab_factor (alpha0 + alpha1) / h.  xsign = None: uncut element (no x-dofs, block nd x nd)."""
    J = (pmat[:dim] - pmat[dim]).T
    meas = abs(np.linalg.det(J))
    h = meas ** (1.0 / dim)
    gphi = np.linalg.inv(J).T @ (ls[:dim] - ls[dim])
    nrm = gphi / np.linalg.norm(gphi)
    nd = orc.load().orc_h1_ndof(dim, order)
    xs = [-1] * nd if xsign is None else xsign
    A = np.zeros((2 * nd, 2 * nd))
    kappa = (kappa_neg, 1.0 - kappa_neg)
    for dt in (NEG, POS):
        npt, pts, w = orc.cut_rule(dim, pmat, ls, dt, 2 * order - 2)
        for q in range(max(npt, 0)):
            U, G = _bvectors(dim, order, vnums, pmat, pts[q], xs, nrm)
            A += meas * w[q] * alpha[dt] * (G[dt] @ G[dt].T)
    npt, pts, w = orc.cut_rule(dim, pmat, ls, IF, 2 * order)
    stab = stab_factor * (alpha[0] + alpha[1]) / h
    for q in range(max(npt, 0)):
        U, G = _bvectors(dim, order, vnums, pmat, pts[q], xs, nrm)
        flux = -sum(kappa[k] * alpha[k] * (G[k] @ nrm) for k in (NEG, POS))
        jump = U[NEG] - U[POS]
        A += meas * w[q] * (np.outer(jump, flux) + np.outer(flux, jump) + stab * np.outer(jump, jump))
    return A if xsign is not None else A[:nd, :nd]


def nxfem_elvec(dim, order, vnums, pmat, ls, xsign, dt, fpoly):
    """compound element vector (2 nd) of SymbolicLFI(levelset_domain = dt, form = f * (v_std + op(v_x))), op = neg for dt = NEG,
    pos for dt = POS (py_tutorials/nxfem.py:109-110; xfem/symboliccutlfi.cpp:52-158, intorder = 2 order :85).
    fpoly: {(ex, ey, ez): coef}.  xsign = None: uncut element (block nd)."""
    J = (pmat[:dim] - pmat[dim]).T
    meas = abs(np.linalg.det(J))
    nd = orc.load().orc_h1_ndof(dim, order)
    xs = [-1] * nd if xsign is None else xsign
    b = np.zeros(2 * nd)
    npt, pts, w = orc.cut_rule(dim, pmat, ls, dt, 2 * order)
    m = (np.asarray(xs) == dt).astype(float)
    for q in range(max(npt, 0)):
        xi = pts[q][:dim]
        s, _ = orc.h1_shape(dim, order, vnums, xi)
        x = np.zeros(3)
        x[:dim] = pmat[dim] + J @ xi
        f = sum(c * x[0] ** e[0] * x[1] ** e[1] * x[2] ** e[2] for e, c in fpoly.items())
        b += meas * w[q] * f * np.concatenate([s, m * s])
    return b if xsign is not None else b[:nd]


# ---------------------------------------------------------------------------------------------------------------------
# the same two functions through the reference's own DiffOpX / XFiniteElement (oracle/_ref, ref_driver.cpp:NxfemElmat):
# what pins the numpy restatement above (tests/test_oracle_vs_ref.py::test_xfem_*)
# ---------------------------------------------------------------------------------------------------------------------
def nxfem_elmat_ref(dim, order, vnums, pmat, ls, xsign, kappa_neg, alpha, stab_factor, dord=0):
    """dord >= 1: pmat = geometry coefficients of the isoparametric element of that order (deformed mesh)"""
    import ctypes
    from . import refload
    L = refload.load()
    nd = orc.load().orc_h1_ndof(dim, order)
    n = 2 * nd if xsign is not None else nd
    em = np.zeros((n, n))
    vn = np.ascontiguousarray(vnums, dtype=np.int32); pm = np.ascontiguousarray(pmat, dtype=np.float64)
    l = np.ascontiguousarray(ls, dtype=np.float64)
    xs = None if xsign is None else np.ascontiguousarray(xsign, dtype=np.int8)
    xp = None if xs is None else xs.ctypes.data_as(ctypes.POINTER(ctypes.c_byte))
    r = L.ref_nxfem_elmat(orc.ET[dim], order, dord, orc.P(vn, orc.ip), orc.P(pm), orc.P(l), xp, kappa_neg, alpha[0], alpha[1], stab_factor, orc.P(em))
    assert r >= 0, r
    return em


def nxfem_elvec_ref(dim, order, vnums, pmat, ls, xsign, dt, fpoly, dord=0):
    import ctypes
    from . import refload
    L = refload.load()
    nd = orc.load().orc_h1_ndof(dim, order)
    n = 2 * nd if xsign is not None else nd
    ev = np.zeros(n)
    vn = np.ascontiguousarray(vnums, dtype=np.int32); pm = np.ascontiguousarray(pmat, dtype=np.float64)
    l = np.ascontiguousarray(ls, dtype=np.float64)
    xs = None if xsign is None else np.ascontiguousarray(xsign, dtype=np.int8)
    xp = None if xs is None else xs.ctypes.data_as(ctypes.POINTER(ctypes.c_byte))
    mono = orc.mono_array([(c, e[0], e[1], e[2]) for e, c in fpoly.items()])
    L.ref_nxfem_elvec(orc.ET[dim], order, dord, orc.P(vn, orc.ip), orc.P(pm), orc.P(l), xp, dt, len(mono), orc.P(mono), orc.P(ev))
    return ev
