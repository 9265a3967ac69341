"""TEST INFRASTRUCTURE -- numpy front-end to the CPU oracle (oracle/libngsx_oracle.so, plain-C
restatement in oracle/ngsx_oracle.c) and to the compiled reference (oracle/_ref, via refload).
Not part of the product path: only tests/, __graft_entry__.smoke() and bench.py's CPU legs import it."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
NEG, POS, IF = 0, 1, 2
ET = {1: 1, 2: 10, 3: 20}   # dim -> NGSolve ELEMENT_TYPE (NGS/fem/elementtopology.hpp:22-25)

c_int, c_i64, c_double = ctypes.c_int, ctypes.c_int64, ctypes.c_double
dp = ctypes.POINTER(ctypes.c_double)
ip = ctypes.POINTER(ctypes.c_int32)
lp = ctypes.POINTER(ctypes.c_int64)
bp = ctypes.POINTER(ctypes.c_int8)
ubp = ctypes.POINTER(ctypes.c_uint8)


def lib_path():
    return os.path.join(_HERE, "libngsx_oracle.so")


def build(force=False):
    if force or not os.path.exists(lib_path()) or \
            os.path.getmtime(lib_path()) < os.path.getmtime(os.path.join(_HERE, "ngsx_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "libngsx_oracle.so"], stdout=subprocess.DEVNULL)


def load():
    global _LIB
    if _LIB is not None:
        return _LIB
    build()
    L = ctypes.CDLL(lib_path())
    L.orc_select_rule.argtypes = [c_int, c_int, dp, dp, c_int]
    L.orc_check_straight_cut.argtypes = [c_int, dp]
    L.orc_cut_rule_untrafo.argtypes = [c_int, dp, c_int, c_int, dp, dp, c_int]
    L.orc_cut_rule.argtypes = [c_int, dp, dp, c_int, c_int, dp, dp, c_int]
    L.orc_cutinfo_element.argtypes = [c_int, dp, dp, dp]
    L.orc_st_cut_rule.argtypes = [c_int, dp, dp, c_int, c_int, c_int, c_int, dp, dp, dp, c_int]
    L.orc_st_integrate_mesh.argtypes = [c_int, c_i64, ip, dp, c_i64, dp, c_int, c_int, c_int, c_int, c_int, dp, lp]
    L.orc_st_integrate_mesh.restype = c_double
    L.orc_st_elmat_h1.argtypes = [c_int, c_int, c_int, dp, dp, c_int, c_int, c_int, c_int, c_double, c_double, c_double, dp]
    L.orc_st_cutinfo_element.argtypes = [c_int, dp, dp, c_int, c_int, dp]
    L.orc_h1_ndof.argtypes = [c_int, c_int]
    L.orc_h1_shape.argtypes = [c_int, c_int, ip, dp, dp, dp]
    L.orc_elmat_h1.argtypes = [c_int, c_int, ip, dp, dp, c_int, c_int, c_double, c_double, dp]
    L.orc_elvec_h1.argtypes = [c_int, c_int, ip, dp, dp, c_int, c_int, c_int, dp, dp]
    L.orc_elmat_h1_cf.argtypes = [c_int, c_int, ip, dp, dp, c_int, c_int, c_double, c_double, c_int, dp, dp]
    L.orc_elvec_h1_cf.argtypes = [c_int, c_int, ip, dp, dp, c_int, c_int, c_int, dp, c_int, dp, dp]
    L.orc_interpolate_p1_clamp.argtypes = [c_i64, dp, c_double, dp]
    L.orc_cutinfo_mesh.argtypes = [c_int, c_i64, ip, dp, dp, bp, dp]
    L.orc_cutinfo_mesh.restype = c_i64
    L.orc_rule_count_mesh.argtypes = [c_int, c_i64, ip, dp, c_int, c_int, ip]
    L.orc_integrate_mesh.argtypes = [c_int, c_i64, ip, dp, dp, c_int, c_int, c_int, dp, lp]
    L.orc_integrate_mesh.restype = c_double
    L.orc_elmat_mesh.argtypes = [c_int, c_i64, lp, ip, dp, dp, c_int, c_int, c_int, c_double, c_double, dp, ip]
    L.orc_elmat_mesh.restype = c_i64
    L.orc_assemble_csr.argtypes = [c_int, ip, dp, dp, c_int, ip, c_int, c_int, c_double, c_double,
                                   c_int, lp, ip, lp, ip, dp]
    L.orc_assemble_csr.restype = c_i64
    L.orc_assemble_vec.argtypes = [c_int, c_i64, ip, dp, dp, c_int, ip, c_int, c_int, c_int, dp, dp]
    L.orc_assemble_vec.restype = c_i64
    L.orc_facets_neighbor_types.argtypes = [c_i64, ip, ubp, ubp, c_int, c_int, c_int, ubp]
    L.orc_elements_with_neighbor_facets.argtypes = [c_i64, c_i64, ip, ubp, ubp]
    L.orc_dofs_of_elements.argtypes = [c_i64, c_int, c_i64, ip, ubp, ubp]
    L.orc_facetpatch_elmat_h1.argtypes = [c_int, c_int, ip, dp, ip, dp, c_double, c_int, dp]
    L.orc_cut_rule_iso.argtypes = [c_int, c_int, dp, dp, c_int, c_int, dp, dp, dp, dp, c_int]
    L.orc_elmat_h1_iso.argtypes = [c_int, c_int, c_int, dp, dp, c_int, c_int, dp, dp]
    L.orc_elvec_h1_iso.argtypes = [c_int, c_int, c_int, dp, dp, c_int, c_int, c_int, dp, dp]
    L.orc_elvec_mesh_iso.argtypes = [c_int, c_i64, lp, ip, dp, dp, c_int, ip, dp, c_int, c_int, c_int, c_int, dp, dp, ip]
    L.orc_elvec_mesh_iso.restype = c_i64
    L.orc_integrate_mesh_iso.argtypes = [c_int, c_i64, ip, dp, dp, c_int, ip, dp, c_int, c_int, c_int, dp, lp]
    L.orc_integrate_mesh_iso.restype = c_double
    L.orc_elmat_mesh_iso.argtypes = [c_int, c_i64, lp, ip, dp, dp, c_int, ip, dp, c_int, c_int, c_int, dp, dp, ip]
    L.orc_elmat_mesh_iso.restype = c_i64
    L.orc_quad_rule_untrafo.argtypes = [dp, c_int, c_int, dp, dp, c_int]
    L.orc_set_threads.argtypes = [c_int]
    L.orc_facetpatch_elmat_h1_iso.argtypes = [c_int, c_int, c_int, ip, dp, ip, dp, c_double, c_int, dp]
    L.orc_csr_pattern.argtypes = [c_i64, c_int, ip, c_i64, lp, ip]
    L.orc_csr_pattern.restype = c_i64
    L.orc_color_elements.argtypes = [c_i64, c_int, ip, c_i64, lp, lp]
    _LIB = L
    return L


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def P(a, t=dp):
    return a.ctypes.data_as(t)


def mono_array(mono):
    """integrand sum_m c * x^a y^b z^c  given as list of (c, a, b, c); None -> constant 1"""
    if mono is None:
        mono = [(1.0, 0, 0, 0)]
    return _d(np.array(mono, dtype=np.float64).reshape(-1, 4))


# ----------------------------------------------------------------------------- single-element wrappers
def cut_rule_untrafo(dim, lset, dt, order, impl="oracle", cap=2048):
    pts = np.zeros((cap, 3)); w = np.zeros(cap)
    ls = _d(lset)
    if impl == "oracle":
        n = load().orc_cut_rule_untrafo(dim, P(ls), dt, order, P(pts), P(w), cap)
    else:
        from . import refload
        n = refload.load().ref_cut_rule_untrafo(ET[dim], P(ls), dt, order, P(pts), P(w), cap)
    if n < 0:
        return n, None, None
    return n, pts[:n].copy(), w[:n].copy()


def cut_rule(dim, pmat, lset, dt, order, impl="oracle", cap=2048):
    pts = np.zeros((cap, 3)); w = np.zeros(cap)
    ls = _d(lset); pm = _d(pmat)
    if impl == "oracle":
        n = load().orc_cut_rule(dim, P(pm), P(ls), dt, order, P(pts), P(w), cap)
    else:
        from . import refload
        n = refload.load().ref_cut_rule(ET[dim], P(pm), P(ls), dt, order, P(pts), P(w), cap)
    if n < 0:
        return n, None, None
    return n, pts[:n].copy(), w[:n].copy()


def st_cut_rule(dim, pmat, lset_st, dt, order_time, order_space, impl="oracle", cap=4096):
    """space-time cut rule of one element: lset_st [ntd][dim+1].  returns (n, pts, tref, weights)"""
    pts = np.zeros((cap, 3)); tq = np.zeros(cap); w = np.zeros(cap)
    ls = _d(lset_st); pm = _d(pmat)
    ntd = ls.shape[0]
    if impl == "oracle":
        n = load().orc_st_cut_rule(dim, P(pm), P(ls), ntd, dt, order_time, order_space, P(pts), P(tq), P(w), cap)
    else:
        from . import refload
        n = refload.load().ref_st_cut_rule(ET[dim], P(pm), P(ls), ntd, dt, order_time, order_space, P(pts), P(tq), P(w), cap)
    if n < 0:
        return n, None, None, None
    return n, pts[:n].copy(), tq[:n].copy(), w[:n].copy()


def st_elmat(dim, pmat, lset_st, fe_order, ntd_fe, dt, order_time, order_space, c_mass, c_stiff, c_dt, impl="oracle"):
    """space-time element matrix of SpaceTimeFE(H1 order fe_order, NodalTimeFE order ntd_fe - 1): (npts, [N][N]), N = ntd_fe * nd_space"""
    ls = _d(lset_st); pm = _d(pmat)
    N = ntd_fe * load().orc_h1_ndof(dim, fe_order)
    out = np.zeros((N, N))
    if impl == "oracle":
        n = load().orc_st_elmat_h1(dim, fe_order, ntd_fe, P(pm), P(ls), ls.shape[0], dt, order_time, order_space, c_mass, c_stiff, c_dt, P(out))
    else:
        from . import refload
        n = refload.load().ref_st_elmat_h1(ET[dim], fe_order, ntd_fe, P(pm), P(ls), ls.shape[0], dt, order_time, order_space,
                                           c_double(c_mass), c_double(c_stiff), c_double(c_dt), P(out))
    return n, out


def st_cutinfo_element(dim, pmat, lset_st, time_order):
    r = c_double(0)
    ls = _d(lset_st); pm = _d(pmat)
    d = load().orc_st_cutinfo_element(dim, P(pm), P(ls), ls.shape[0], time_order, ctypes.byref(r))
    return d, r.value


def st_integrate_mesh(dim, el2v, vert, lset_st, dt, order_space, order_time, mono5=None, nthreads=0):
    """IntegrateX(..., time_order=order_time): lset_st [ntd][nv]; mono5 rows (coef, ex, ey, ez, et)"""
    el2v = _i(el2v); vert = _d(vert); ls = _d(lset_st)
    mo = _d(np.array([[1.0, 0, 0, 0, 0]] if mono5 is None else mono5, dtype=np.float64).reshape(-1, 5))
    npt = c_i64(0)
    load().orc_set_threads(nthreads)
    v = load().orc_st_integrate_mesh(dim, el2v.shape[0], P(el2v, ip), P(vert), vert.shape[0], P(ls), ls.shape[0], dt, order_space, order_time,
                                     mo.shape[0], P(mo), ctypes.byref(npt))
    return v, npt.value


def cutinfo_element(dim, pmat, lset, impl="oracle"):
    r = c_double(0)
    ls = _d(lset); pm = _d(pmat)
    if impl == "oracle":
        d = load().orc_cutinfo_element(dim, P(pm), P(ls), ctypes.byref(r))
    else:
        from . import refload
        d = refload.load().ref_cutinfo_element(ET[dim], P(pm), P(ls), ctypes.byref(r))
    return d, r.value


def h1_shape(dim, order, vnums, xi, impl="oracle"):
    nd = load().orc_h1_ndof(dim, order)
    sh = np.zeros(nd); dsh = np.zeros((nd, dim))
    x3 = np.zeros(3); x3[:len(xi)] = xi
    vn = _i(vnums) if vnums is not None else None
    vp = P(vn, ip) if vn is not None else None
    if impl == "oracle":
        n = load().orc_h1_shape(dim, order, vp, P(x3), P(sh), P(dsh))
    else:
        from . import refload
        n = refload.load().ref_h1_shape(ET[dim], order, vp, P(x3), P(sh), P(dsh))
    assert n == nd, (n, nd)
    return sh, dsh


def elmat_h1(dim, fe_order, vnums, pmat, lset, dt, intorder, c_mass, c_stiff, impl="oracle"):
    nd = load().orc_h1_ndof(dim, fe_order)
    em = np.zeros((nd, nd))
    vn = _i(vnums); pm = _d(pmat); ls = _d(lset)
    if impl == "oracle":
        n = load().orc_elmat_h1(dim, fe_order, P(vn, ip), P(pm), P(ls), dt, intorder, c_mass, c_stiff, P(em))
    else:
        from . import refload
        n = refload.load().ref_elmat_h1(ET[dim], fe_order, P(vn, ip), P(pm), P(ls), dt, intorder, c_mass, c_stiff, P(em))
    return n, em


def elvec_h1(dim, fe_order, vnums, pmat, lset, dt, intorder, mono, impl="oracle"):
    nd = load().orc_h1_ndof(dim, fe_order)
    ev = np.zeros(nd)
    vn = _i(vnums); pm = _d(pmat); ls = _d(lset); mo = mono_array(mono)
    if impl == "oracle":
        n = load().orc_elvec_h1(dim, fe_order, P(vn, ip), P(pm), P(ls), dt, intorder, len(mo), P(mo), P(ev))
    else:
        from . import refload
        n = refload.load().ref_elvec_h1(ET[dim], fe_order, P(vn, ip), P(pm), P(ls), dt, intorder, len(mo), P(mo), P(ev))
    return n, ev


def elmat_h1_cf(dim, fe_order, vnums, pmat, lset, dt, intorder, c_mass, c_stiff, cf_order, cf_local):
    """element matrix of gf * (c_stiff grad(u) grad(v) + c_mass u v): cf_local = the coefficient GridFunction on the element's dofs"""
    nd = load().orc_h1_ndof(dim, fe_order)
    em = np.zeros((nd, nd))
    vn = _i(vnums); pm = _d(pmat); ls = _d(lset); cl = _d(cf_local)
    n = load().orc_elmat_h1_cf(dim, fe_order, P(vn, ip), P(pm), P(ls), dt, intorder, c_mass, c_stiff, cf_order, P(cl), P(em))
    return n, em


def elvec_h1_cf(dim, fe_order, vnums, pmat, lset, dt, intorder, mono, cf_order, cf_local):
    """element vector of (poly * gf) v: cf_local = the coefficient GridFunction's values on the element's dofs (H1 order cf_order)"""
    nd = load().orc_h1_ndof(dim, fe_order)
    ev = np.zeros(nd)
    vn = _i(vnums); pm = _d(pmat); ls = _d(lset); mo = mono_array(mono); cl = _d(cf_local)
    n = load().orc_elvec_h1_cf(dim, fe_order, P(vn, ip), P(pm), P(ls), dt, intorder, len(mo), P(mo), cf_order, P(cl), P(ev))
    return n, ev


def facetpatch_elmat_h1(dim, fe_order, vn1, pmat1, vn2, pmat2, coef, hpow=0, impl="oracle", dord=0):
    """impl="ref": the compiled reference (MapPatchIntegrationPoint of xfem/symboliccutbfi.cpp + H1HighOrderFE); dord >= 1:
    pmat1 / pmat2 are the coefficients of the isoparametric geometry elements (deformed mesh)"""
    nd = load().orc_h1_ndof(dim, fe_order)
    em = np.zeros((2 * nd, 2 * nd))
    a, b, p1, p2 = _i(vn1), _i(vn2), _d(pmat1), _d(pmat2)
    if impl == "oracle":
        if dord == 0:
            n = load().orc_facetpatch_elmat_h1(dim, fe_order, P(a, ip), P(p1), P(b, ip), P(p2), coef, hpow, P(em))
        else:
            n = load().orc_facetpatch_elmat_h1_iso(dim, fe_order, dord, P(a, ip), P(p1), P(b, ip), P(p2), coef, hpow, P(em))
    else:
        from . import refload
        n = refload.load().ref_facetpatch_elmat_h1(ET[dim], fe_order, dord, P(a, ip), P(p1), P(b, ip), P(p2), coef, hpow, P(em), None)
    return n, em


def elmat_h1_form(dim, fe_order, vnums, geom, lset, dt, intorder, form, dord=0):
    """compiled reference only: general ngsx_form_h1 element matrix (Nitsche / dn.dn terms, mesh_size powers), affine (dord 0:
    geom = vertex matrix) or isoparametric (geom = geometry coefficients)"""
    from . import refload
    nd = load().orc_h1_ndof(dim, fe_order)
    em = np.zeros((nd, nd))
    vn = _i(vnums) if vnums is not None else None
    g = _d(geom); ls = _d(lset); f = _d(form)
    n = refload.load().ref_elmat_h1_form(ET[dim], fe_order, dord, P(vn, ip) if vn is not None else None, P(g), P(ls), dt, intorder, P(f), P(em))
    return n, em


# ----------------------------------------------------------------------------- mesh-level wrappers
def cutinfo_mesh(dim, el2v, vert, lset_v, impl="oracle", nthreads=0):
    el2v = _i(el2v); vert = _d(vert); lset_v = _d(lset_v)
    ne = el2v.shape[0]
    dom = np.zeros(ne, dtype=np.int8); ratio = np.zeros(ne)
    if impl == "oracle":
        load().orc_set_threads(nthreads)
        nerr = load().orc_cutinfo_mesh(dim, ne, P(el2v, ip), P(vert), P(lset_v), P(dom, bp), P(ratio))
    else:
        from . import refload
        nerr = refload.load().ref_cutinfo_mesh(dim, ne, P(el2v, ip), P(vert), P(lset_v),
                                               dom.ctypes.data_as(ctypes.POINTER(ctypes.c_byte)), P(ratio), nthreads)
    return dom, ratio, nerr


def rule_count_mesh(dim, el2v, lset_v, dt, order):
    el2v = _i(el2v); lset_v = _d(lset_v)
    npts = np.zeros(el2v.shape[0], dtype=np.int32)
    load().orc_rule_count_mesh(dim, el2v.shape[0], P(el2v, ip), P(lset_v), dt, order, P(npts, ip))
    return npts


def integrate_mesh(dim, el2v, vert, lset_v, dt, order, mono=None, impl="oracle", nthreads=0):
    el2v = _i(el2v); vert = _d(vert); lset_v = _d(lset_v); mo = mono_array(mono)
    npt = c_i64(0)
    if impl == "oracle":
        load().orc_set_threads(nthreads)
        v = load().orc_integrate_mesh(dim, el2v.shape[0], P(el2v, ip), P(vert), P(lset_v), dt, order, len(mo), P(mo),
                                      ctypes.byref(npt))
    else:
        from . import refload
        npl = ctypes.c_long(0)
        v = refload.load().ref_integrate_mesh(dim, el2v.shape[0], P(el2v, ip), P(vert), P(lset_v), dt, order, len(mo), P(mo),
                                              nthreads, ctypes.byref(npl))
        npt = c_i64(npl.value)
    return v, npt.value


def elmat_mesh(dim, el2v, vert, lset_v, fe_order, dt, intorder, c_mass, c_stiff, els=None, impl="oracle", nthreads=0):
    el2v = _i(el2v); vert = _d(vert); lset_v = _d(lset_v)
    nd = load().orc_h1_ndof(dim, fe_order)
    if els is not None:
        els = np.ascontiguousarray(els, dtype=np.int64)
    nel = el2v.shape[0] if els is None else len(els)
    em = np.zeros((nel, nd, nd)); npts = np.zeros(nel, dtype=np.int32)
    elp = P(els, lp) if els is not None else None
    if impl == "oracle":
        load().orc_set_threads(nthreads)
        load().orc_elmat_mesh(dim, nel, elp, P(el2v, ip), P(vert), P(lset_v), fe_order, dt, intorder, c_mass, c_stiff,
                              P(em), P(npts, ip))
    else:
        from . import refload
        elp = els.ctypes.data_as(ctypes.POINTER(ctypes.c_long)) if els is not None else None
        refload.load().ref_elmat_mesh(dim, nel, elp, P(el2v, ip), P(vert), P(lset_v), fe_order, dt, intorder, c_mass, c_stiff,
                                      P(em), P(npts, ip), nthreads)
    return em, npts


def assemble_csr(dim, el2v, vert, lset_v, fe_order, el2dof, dt, intorder, c_mass, c_stiff,
                 color_ptr, color_els, rowptr, colidx, impl="oracle", nthreads=0):
    el2v = _i(el2v); vert = _d(vert); lset_v = _d(lset_v); el2dof = _i(el2dof)
    color_ptr = np.ascontiguousarray(color_ptr, dtype=np.int64); color_els = _i(color_els)
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64); colidx = _i(colidx)
    vals = np.zeros(len(colidx))
    if impl == "oracle":
        load().orc_set_threads(nthreads)
        n = load().orc_assemble_csr(dim, P(el2v, ip), P(vert), P(lset_v), fe_order, P(el2dof, ip), dt, intorder, c_mass, c_stiff,
                                    len(color_ptr) - 1, P(color_ptr, lp), P(color_els, ip), P(rowptr, lp), P(colidx, ip), P(vals))
    else:
        from . import refload
        clp = ctypes.POINTER(ctypes.c_long)
        n = refload.load().ref_assemble_csr(dim, P(el2v, ip), P(vert), P(lset_v), fe_order, P(el2dof, ip), dt, intorder,
                                            c_mass, c_stiff, len(color_ptr) - 1, color_ptr.ctypes.data_as(clp), P(color_els, ip),
                                            rowptr.ctypes.data_as(clp), P(colidx, ip), P(vals), nthreads)
    return vals, n


def assemble_vec(dim, el2v, vert, lset_v, fe_order, el2dof, ndof, dt, intorder, mono):
    el2v = _i(el2v); vert = _d(vert); lset_v = _d(lset_v); el2dof = _i(el2dof); mo = mono_array(mono)
    vec = np.zeros(ndof)
    load().orc_assemble_vec(dim, el2v.shape[0], P(el2v, ip), P(vert), P(lset_v), fe_order, P(el2dof, ip), dt, intorder,
                            len(mo), P(mo), P(vec))
    return vec


def facets_neighbor_types(facet2el, a, b, bnd_a=True, bnd_b=True, use_and=True):
    f2e = _i(facet2el); a = np.ascontiguousarray(a, dtype=np.uint8); b = np.ascontiguousarray(b, dtype=np.uint8)
    out = np.zeros(f2e.shape[0], dtype=np.uint8)
    load().orc_facets_neighbor_types(f2e.shape[0], P(f2e, ip), P(a, ubp), P(b, ubp), int(bnd_a), int(bnd_b), int(use_and), P(out, ubp))
    return out.astype(bool)


def elements_with_neighbor_facets(facet2el, ne, a):
    f2e = _i(facet2el); a = np.ascontiguousarray(a, dtype=np.uint8)
    out = np.zeros(ne, dtype=np.uint8)
    load().orc_elements_with_neighbor_facets(f2e.shape[0], ne, P(f2e, ip), P(a, ubp), P(out, ubp))
    return out.astype(bool)


def dofs_of_elements(el2dof, ndof, a):
    e2d = _i(el2dof); a = np.ascontiguousarray(a, dtype=np.uint8)
    out = np.zeros(ndof, dtype=np.uint8)
    load().orc_dofs_of_elements(e2d.shape[0], e2d.shape[1], ndof, P(e2d, ip), P(a, ubp), P(out, ubp))
    return out.astype(bool)


# ----------------------------------------------------------------------------- deformed (isoparametric) elements, f-4
def form8(c_mass=0.0, c_stiff=0.0, c_nitsche=0.0, c_dndn=0.0, hpow=(0, 0, 0, 0)):
    return _d([c_mass, c_stiff, c_nitsche, c_dndn] + [float(p) for p in hpow])


def cut_rule_iso(dim, dord, gcoef, lset, dt, order, impl="oracle", cap=2048):
    """points, weights (IF: times |F^-T n| per point), measure |det F| and mapped point per rule point"""
    pts = np.zeros((cap, 3)); w = np.zeros(cap); meas = np.zeros(cap); xp = np.zeros((cap, 3))
    gc = _d(gcoef); ls = _d(lset)
    if impl == "oracle":
        n = load().orc_cut_rule_iso(dim, dord, P(gc), P(ls), dt, order, P(pts), P(w), P(meas), P(xp), cap)
    else:
        from . import refload
        n = refload.load().ref_cut_rule_iso(ET[dim], dord, P(gc), P(ls), dt, order, P(pts), P(w), P(meas), P(xp), cap)
    if n < 0:
        return n, None, None, None, None
    return n, pts[:n], w[:n], meas[:n], xp[:n]


def elmat_h1_iso(dim, fe_order, dord, gcoef, lset, dt, intorder, form, impl="oracle"):
    nd = load().orc_h1_ndof(dim, fe_order)
    em = np.zeros((nd, nd))
    gc = _d(gcoef); ls = _d(lset); f = _d(form)
    if impl == "oracle":
        n = load().orc_elmat_h1_iso(dim, fe_order, dord, P(gc), P(ls), dt, intorder, P(f), P(em))
    else:
        from . import refload
        assert f[2] == 0 and f[3] == 0 and not f[4:].any(), "the compiled-reference driver restates mass + stiffness only"
        n = refload.load().ref_elmat_h1_iso(ET[dim], fe_order, dord, P(gc), P(ls), dt, intorder, f[0], f[1], P(em))
    return n, em


def integrate_mesh_iso(dim, el2v, vert, lset_v, dord, el2dof_def, defo, dt, order, mono=None, nthreads=0):
    ev = _i(el2v); vt = _d(vert); ls = _d(lset_v); ed = _i(el2dof_def); df = _d(defo); mo = mono_array(mono)
    load().orc_set_threads(nthreads)
    npt = c_i64(0)
    val = load().orc_integrate_mesh_iso(dim, ev.shape[0], P(ev, ip), P(vt), P(ls), dord, P(ed, ip), P(df), dt, order, len(mo), P(mo),
                                        ctypes.byref(npt))
    return val, npt.value


def elmat_mesh_iso(dim, el2v, vert, lset_v, dord, el2dof_def, defo, fe_order, dt, intorder, form, els=None, nthreads=0):
    ev = _i(el2v); vt = _d(vert); ls = _d(lset_v); ed = _i(el2dof_def); df = _d(defo); f = _d(form)
    nd = load().orc_h1_ndof(dim, fe_order)
    load().orc_set_threads(nthreads)
    if els is None:
        nel, ep = ev.shape[0], None
    else:
        ee = np.ascontiguousarray(els, dtype=np.int64); nel, ep = len(ee), P(ee, lp)
    out = np.zeros((nel, nd, nd)); npts = np.zeros(nel, dtype=np.int32)
    load().orc_elmat_mesh_iso(dim, nel, ep, P(ev, ip), P(vt), P(ls), dord, P(ed, ip), P(df), fe_order, dt, intorder, P(f), P(out), P(npts, ip))
    return out, npts


def elvec_h1_iso(dim, fe_order, dord, gcoef, lset, dt, intorder, mono, impl="oracle"):
    nd = load().orc_h1_ndof(dim, fe_order)
    evv = np.zeros(nd)
    gc = _d(gcoef); ls = _d(lset); mo = mono_array(mono)
    if impl == "oracle":
        n = load().orc_elvec_h1_iso(dim, fe_order, dord, P(gc), P(ls), dt, intorder, len(mo), P(mo), P(evv))
    else:
        from . import refload
        n = refload.load().ref_elvec_h1_iso(ET[dim], fe_order, dord, P(gc), P(ls), dt, intorder, len(mo), P(mo), P(evv))
    return n, evv


def elvec_mesh_iso(dim, el2v, vert, lset_v, dord, el2dof_def, defo, fe_order, dt, intorder, mono, els=None, nthreads=0):
    ev = _i(el2v); vt = _d(vert); ls = _d(lset_v); ed = _i(el2dof_def); df = _d(defo); mo = mono_array(mono)
    nd = load().orc_h1_ndof(dim, fe_order)
    load().orc_set_threads(nthreads)
    if els is None:
        nel, ep = ev.shape[0], None
    else:
        ee = np.ascontiguousarray(els, dtype=np.int64); nel, ep = len(ee), P(ee, lp)
    out = np.zeros((nel, nd)); npts = np.zeros(nel, dtype=np.int32)
    load().orc_elvec_mesh_iso(dim, nel, ep, P(ev, ip), P(vt), P(ls), dord, P(ed, ip), P(df), fe_order, dt, intorder, len(mo), P(mo), P(out), P(npts, ip))
    return out, npts


def quad_rule_untrafo(vals, dt, order, cap=4096):
    """cut rule on the reference quadrilateral (C twin of oracle/quad_ref.py:rule_untrafo)"""
    v = _d(vals); pts = np.zeros((cap, 3)); w = np.zeros(cap)
    n = load().orc_quad_rule_untrafo(P(v), dt, order, P(pts), P(w), cap)
    return n, pts[:max(n, 0)], w[:max(n, 0)]
