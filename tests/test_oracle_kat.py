"""CPU oracle against the reference's OWN known-answer tests for the path:
tests/pytests/test_localgeom.py:24-45, test_straightcutrule.py:11-31, 111-126, 134-154, 166-185, 191-207, 218-235,
242-280, test_xfes_ndof.py:5-21 -- same level sets, same analytic values, same tolerances."""
from math import pi, sqrt

import numpy as np
import pytest

from oracle import orc
from ngsxfem_b200 import meshes

NEG, POS, IF = 0, 1, 2


def p1(levelset, vert, eps=1e-14):
    v = levelset(*vert.T)
    return np.where(np.abs(v) < eps, eps, v)


def integrate(m, ls, dt, order, mono=None):
    return orc.integrate_mesh(m.dim, m.el2v.numpy(), m.vert.numpy(), ls, dt, order, mono)[0]


def test_cut_triangle_localgeom():
    # test_localgeom.py:24-45: triangle (0,0),(1,0),(0,1) in 4 elements, levelset x+y-0.25, orders 0..15
    vert = np.array([[0, 0], [1, 0], [0, 1], [.5, 0], [.5, .5], [0, .5]], dtype=float)
    el2v = np.array([[0, 3, 5], [3, 1, 4], [5, 4, 2], [3, 4, 5]], dtype=np.int32)
    ls = p1(lambda x, y: x + y - 0.25, vert)
    for order in range(16):
        neg = orc.integrate_mesh(2, el2v, vert, ls, NEG, order)[0]
        pos = orc.integrate_mesh(2, el2v, vert, ls, POS, order)[0]
        tol = 5e-16 * (order + 1) ** 2
        assert abs(neg - 1 / 32) < tol
        assert abs(pos - 0.5 + 1 / 32) < tol
        assert abs(neg + pos - 0.5) < tol


@pytest.mark.parametrize("domain", [NEG, POS, IF])
@pytest.mark.parametrize("alpha", [2, 4, 8])
def test_polynomial_segm(domain, alpha):
    # test_straightcutrule.py:11-31 (1D): the oracle's 1D rule on the unit segment, integrand x^alpha
    x_ast = 0.78522
    ref = {POS: x_ast ** (alpha + 1) / (alpha + 1), NEG: (1 - x_ast ** (alpha + 1)) / (alpha + 1), IF: x_ast ** alpha}
    pm = np.array([[0.0], [1.0]])                       # Make1DMesh(1): vertices 0 and 1
    ls = np.array([x_ast - 0.0, x_ast - 1.0])
    n, pts, w = orc.cut_rule(1, pm, ls, domain, alpha)
    xs = pm[1, 0] + (pm[0, 0] - pm[1, 0]) * pts[:, 0]
    val = float(np.sum(w * xs ** alpha)) * abs(pm[0, 0] - pm[1, 0])
    assert abs(val - ref[domain]) < 5e-15 * (alpha + 1) ** 2


@pytest.mark.parametrize("order", [2, 4, 8])
@pytest.mark.parametrize("domain", [NEG, POS, IF])
@pytest.mark.parametrize("N", [1, 10, 30])
def test_straight_cut_2d(order, domain, N):
    # test_straightcutrule.py:111-126 (trig branch)
    m = meshes.MakeStructured2DMesh(N)
    ls = p1(lambda x, y: 1 - 2 * x - 2 * y, m.vert.numpy())
    ref = {NEG: 7 / 8, POS: 1 / 8, IF: 1 / sqrt(2)}
    assert abs(integrate(m, ls, domain, order) - ref[domain]) < 5e-15 * (order + 1) ** 2


@pytest.mark.parametrize("domain", [IF, NEG, POS])
@pytest.mark.parametrize("dim", [0, 1])
@pytest.mark.parametrize("eps", [1e-1, 1e-2, 5e-3, 1e-3, 0])
def test_orth_cut_2d_epsiloned(domain, dim, eps):
    # test_straightcutrule.py:134-154
    m = meshes.MakeStructured2DMesh(1)
    f = (lambda x, y: 1 - 2 * x + eps * (y - 0.5)) if dim == 0 else (lambda x, y: 1 - 2 * y + eps * (x - 0.5))
    ls = p1(f, m.vert.numpy())
    ref = {NEG: 0.5, POS: 0.5, IF: sqrt(1 + eps * eps / 4)}
    assert abs(integrate(m, ls, domain, 2) - ref[domain]) < 5e-15 * 9


@pytest.mark.parametrize("order", [2, 4, 8])
@pytest.mark.parametrize("domain", [NEG, POS])
@pytest.mark.parametrize("alpha", [0, 1, 2])
@pytest.mark.parametrize("dim", [0, 1])
def test_straight_cut_2d_polynomial(order, domain, alpha, dim):
    # test_straightcutrule.py:166-185
    m = meshes.MakeStructured2DMesh(1)
    ls = p1(lambda x, y: 1 - 2 * x - 2 * y, m.vert.numpy())
    val_pos = 2 ** (-alpha - 2) / (alpha * alpha + 3 * alpha + 2)
    ref = {POS: val_pos, NEG: 1 / (alpha + 1) - val_pos}
    mono = [(1.0, alpha if dim == 0 else 0, alpha if dim == 1 else 0, 0)]
    assert abs(integrate(m, ls, domain, order, mono) - ref[domain]) < 5e-15 * (order + 1) ** 2


@pytest.mark.parametrize("order", [2, 4, 8])
@pytest.mark.parametrize("domain", [NEG, POS, IF])
def test_straight_cut_3d(order, domain):
    # test_straightcutrule.py:191-207 (tet branch)
    m = meshes.MakeStructured3DMesh(1)
    ls = p1(lambda x, y, z: 1 - 2 * x - 2 * y - 2 * z, m.vert.numpy())
    ref = {POS: 1 / 48, NEG: 47 / 48, IF: sqrt(3) / 8}
    assert abs(integrate(m, ls, domain, order) - ref[domain]) < 5e-15 * (order + 1) ** 2


@pytest.mark.parametrize("domain", [NEG, POS])
@pytest.mark.parametrize("alpha", [0, 1, 2])
@pytest.mark.parametrize("dim", [0, 1, 2])
def test_straight_cut_3d_polynomial(domain, alpha, dim):
    # test_straightcutrule.py:218-235
    order = 4
    m = meshes.MakeStructured3DMesh(1)
    ls = p1(lambda x, y, z: 1 - 2 * x - 2 * y - 2 * z, m.vert.numpy())
    val_pos = 2 ** (-alpha - 3) / (alpha ** 3 + 6 * alpha * alpha + 11 * alpha + 6)
    ref = {POS: val_pos, NEG: 1 / (alpha + 1) - val_pos}
    e = [0, 0, 0]
    e[dim] = alpha
    assert abs(integrate(m, ls, domain, order, [(1.0, *e)]) - ref[domain]) < 5e-15 * (order + 1) ** 2


@pytest.mark.parametrize("domain", [NEG, POS, IF])
@pytest.mark.parametrize("dim", [0, 1])
def test_orth_cut_2d(domain, dim):
    # test_straightcutrule.py:242-257
    m = meshes.MakeStructured2DMesh(1)
    ls = p1((lambda x, y: 1 - 3 * x) if dim == 0 else (lambda x, y: 1 - 3 * y), m.vert.numpy())
    ref = {NEG: 2 / 3, POS: 1 / 3, IF: 1.0}
    assert abs(integrate(m, ls, domain, 2) - ref[domain]) < 5e-15 * 9


@pytest.mark.parametrize("domain", [NEG, POS])
@pytest.mark.parametrize("dim", [0, 1, 2])
def test_orth_cut_3d(domain, dim):
    # test_straightcutrule.py:265-280
    m = meshes.MakeStructured3DMesh(1)
    fs = [lambda x, y, z: 1 - 2 * x, lambda x, y, z: 1 - 2 * y, lambda x, y, z: 1 - 2 * z]
    ls = p1(fs[dim], m.vert.numpy())
    assert abs(integrate(m, ls, domain, 2) - 0.5) < 5e-15 * 9


@pytest.mark.parametrize("order", [2, 4, 8])
@pytest.mark.parametrize("domain", [NEG, POS, IF])
def test_circle_eoc(order, domain):
    # test_straightcutrule.py:37-68 (trig branch): second-order convergence of the P1 geometry
    r = 0.6
    ref = {POS: 1 - pi * r * r / 4, NEG: pi * r * r / 4, IF: r * pi / 2}
    errs = []
    for i in range(8):
        m = meshes.MakeStructured2DMesh(2 ** i)
        ls = p1(lambda x, y: np.sqrt(x * x + y * y) - r, m.vert.numpy())
        errs.append(abs(integrate(m, ls, domain, order) - ref[domain]))
    eoc = [np.log(errs[i + 1] / errs[i]) / np.log(0.5) for i in range(7)]
    assert np.mean(eoc[1:]) > 1.75


def test_xfes_ndof():
    # test_xfes_ndof.py:5-21: XFESpace.ndof = #P1 dofs of IF elements = 7 (2D 4x4) / 35 (3D 4^3, cylinder level set)
    m = meshes.MakeStructured2DMesh(4, lo=(-1, -1), hi=(1, 1))
    ls = p1(lambda x, y: np.sqrt(x * x + y * y) - 1 / 3, m.vert.numpy())
    dom, _, nerr = orc.cutinfo_mesh(2, m.el2v.numpy(), m.vert.numpy(), ls)
    assert nerr == 0 and m.nv == 25
    assert orc.dofs_of_elements(m.el2v.numpy(), m.nv, dom == IF).sum() == 7
    m = meshes.MakeStructured3DMesh(4, lo=(-1, -1, -1), hi=(1, 1, 1))
    ls = p1(lambda x, y, z: np.sqrt(x * x + y * y) - 1 / 3, m.vert.numpy())
    dom, _, nerr = orc.cutinfo_mesh(3, m.el2v.numpy(), m.vert.numpy(), ls)
    assert nerr == 0 and m.nv == 125
    assert orc.dofs_of_elements(m.el2v.numpy(), m.nv, dom == IF).sum() == 35


# ---- deformed (isoparametric) meshes -------------------------------------------------------------
@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("dord", [1, 2])
def test_deformed_similarity_and_zero_deformation(dim, dord):
    """closed forms for the oracle's isoparametric loops: a zero deformation reproduces the affine integrals; the similarity
    d(x) = (s R - I) x scales volumes by s^dim and the interface by s^(dim-1) (per-point |F^-T n|, cutint/straightcutrule.cpp:558-576)"""
    m = meshes.MakeStructured2DMesh(8) if dim == 2 else meshes.MakeStructured3DMesh(4)
    vert, el2v = m.vert.numpy(), m.el2v.numpy()
    ls = p1((lambda x, y: np.sqrt((x - .5) ** 2 + (y - .5) ** 2) - 0.33) if dim == 2 else
            (lambda x, y, z: np.sqrt((x - .5) ** 2 + (y - .5) ** 2 + (z - .5) ** 2) - 0.33), vert)
    e2d, ndof = m.h1_dofs(dord)
    e2d = e2d.numpy()
    s, th = 1.7, 0.4
    R = np.eye(dim)
    R[0, 0], R[0, 1], R[1, 0], R[1, 1] = np.cos(th), -np.sin(th), np.sin(th), np.cos(th)
    zero = np.zeros((ndof, dim))
    sim = zero.copy()
    sim[:m.nv] = vert @ (s * R - np.eye(dim)).T
    mono = [(1.0, 0, 0, 0)]
    for dt in (NEG, POS, IF):
        base = integrate(m, ls, dt, 2)
        z = orc.integrate_mesh_iso(dim, el2v, vert, ls, dord, e2d, zero, dt, 2, mono)[0]
        g = orc.integrate_mesh_iso(dim, el2v, vert, ls, dord, e2d, sim, dt, 2, mono)[0]
        assert abs(z - base) <= 1e-13 * abs(base)
        assert abs(g - base * (s ** dim if dt != IF else s ** (dim - 1))) <= 1e-12 * abs(g)
    # a curved (order-2) boundary-preserving deformation keeps NEG + POS = |Omega| = 1
    if dord == 2:
        rng = np.random.default_rng(5)
        bub = zero.copy()
        e2v, _ = m.edges()
        e2v = e2v.numpy()
        on_bnd = lambda p: np.any((np.abs(p) < 1e-12) | (np.abs(p - 1) < 1e-12), axis=1)
        inner_edge = ~(on_bnd(vert[e2v[:, 0]]) & on_bnd(vert[e2v[:, 1]]))
        bub[m.nv:][inner_edge] = 0.05 * rng.uniform(-1, 1, (int(inner_edge.sum()), dim))
        tot = sum(orc.integrate_mesh_iso(dim, el2v, vert, ls, dord, e2d, bub, dt, 4, mono)[0] for dt in (NEG, POS))
        assert abs(tot - 1.0) <= 1e-12


# ---- SURVEY section 8 row f-5 groundwork: the reference's quad / hex known answers on the restatement oracle/quad_ref.py ----------
QUAD_V = [(0, 0), (1, 0), (1, 1), (0, 1)]
HEX_V = [(0, 0, 0), (1, 0, 0), (1, 1, 0), (0, 1, 0), (0, 0, 1), (1, 0, 1), (1, 1, 1), (0, 1, 1)]


@pytest.mark.parametrize("domain", [IF, NEG, POS])
@pytest.mark.parametrize("dim", [0, 1])
@pytest.mark.parametrize("eps", [1e-1, 1e-2, 5e-3, 1e-3, 0])
def test_orth_cut_quad2d_epsiloned(domain, dim, eps):
    # test_straightcutrule.py:128-157 with quads = True: one unit square, levelset 1 - 2 x + eps (y - 1/2) (or x <-> y), order 2
    from oracle import quad_ref as qr
    f = (lambda x, y: 1 - 2 * x + eps * (y - 0.5)) if dim == 0 else (lambda x, y: 1 - 2 * y + eps * (x - 0.5))
    vals = p1(f, np.array(QUAD_V, dtype=float))
    n, _, w = qr.rule_untrafo(vals, domain, 2)
    ref = {NEG: 0.5, POS: 0.5, IF: sqrt(1.0 + eps * eps / 4)}[domain]
    assert n > 0 and abs(w.sum() - ref) < 5e-15 * 9


@pytest.mark.parametrize("domain", [NEG, POS, IF])
@pytest.mark.parametrize("dim", [0, 1])
def test_orth_cut_quad2d(domain, dim):
    # test_straightcutrule.py:242-262 with quads = True: levelset 1 - 3 x (or y)
    from oracle import quad_ref as qr
    vals = p1((lambda x, y: 1 - 3 * x) if dim == 0 else (lambda x, y: 1 - 3 * y), np.array(QUAD_V, dtype=float))
    n, _, w = qr.rule_untrafo(vals, domain, 2)
    assert n > 0 and abs(w.sum() - {NEG: 2 / 3, POS: 1 / 3, IF: 1.0}[domain]) < 5e-15 * 9


@pytest.mark.parametrize("domain", [NEG, POS, IF])
@pytest.mark.parametrize("dim", [0, 1, 2])
def test_orth_cut_hex3d(domain, dim):
    # test_straightcutrule.py:265-285 with hexes = True: levelset 1 - 2 x (y, z) on one unit cube
    from oracle import quad_ref as qr
    vals = p1(lambda x, y, z: 1 - 2 * (x, y, z)[dim], np.array(HEX_V, dtype=float))
    n, _, w = qr.rule_untrafo_hex(vals, domain, 2)
    assert n > 0 and abs(w.sum() - {NEG: 0.5, POS: 0.5, IF: 1.0}[domain]) < 5e-15 * 9


@pytest.mark.parametrize("dim", [2, 3])
def test_elvec_with_gridfunction_coefficient_is_exact_for_polynomials(dim):
    """orc_elvec_h1_cf (f = poly * GridFunction, the GridFunction evaluated as sum_i c_i phi_i like NGSolve's
    GridFunctionCoefficientFunction): a P2 GridFunction that holds a quadratic polynomial exactly -- vertex values + hierarchical
    edge coefficients c_e = -8 (f(mid) - (f(a) + f(b)) / 2), since phi_e(mid) = -1/8 (NGS/fem/recursive_pol.hpp:1372) -- reproduces the
    element vector of the polynomial coefficient (itself pinned against the compiled reference) on cut and uncut elements"""
    rng = np.random.default_rng(31 + dim)
    edges = [(2, 0), (1, 2), (0, 1)] if dim == 2 else [(3, 0), (3, 1), (3, 2), (0, 1), (0, 2), (1, 2)]
    f = (lambda X: 1.0 + X[..., 0] * X[..., 1] + 0.5 * X[..., -1] ** 2 - 0.3 * X[..., 0])
    mono = [(1.0, 0, 0, 0), (1.0, 1, 1, 0), (-0.3, 1, 0, 0)] + ([(0.5, 0, 0, 2)] if dim == 3 else [(0.5, 0, 2, 0)])
    nchk = 0
    for trial in range(40):
        pm = rng.uniform(-1, 1, size=(dim + 1, dim))
        if abs(np.linalg.det((pm[:dim] - pm[dim]).T)) < 0.05:
            continue
        ls = rng.uniform(-1, 1, size=dim + 1) + (0.0 if trial % 3 else 2.0)
        vn = np.arange(dim + 1, dtype=np.int32)
        fv = f(pm)
        cf2 = np.concatenate([fv, [-8.0 * (f(0.5 * (pm[a] + pm[b])) - 0.5 * (fv[a] + fv[b])) for a, b in edges]])
        for dt in (0, 1, 2):
            for fo in (1, 2):
                n1, ref = orc.elvec_h1(dim, fo, vn, pm, ls, dt, 2 * fo + 2, mono)
                n2, got = orc.elvec_h1_cf(dim, fo, vn, pm, ls, dt, 2 * fo + 2, [(1.0, 0, 0, 0)], 2, cf2)
                assert max(n1, 0) == max(n2, 0)
                if n1 > 0:
                    nchk += 1
                    assert np.abs(got - ref).max() <= 1e-13 * max(np.abs(ref).max(), 1e-3)
                    # a P1 coefficient times a polynomial: (x) * (P1 holding 1 + 2y) against the product polynomial
                    p1 = 1.0 + 2.0 * pm[:, 1]
                    _, a1 = orc.elvec_h1_cf(dim, fo, vn, pm, ls, dt, 2 * fo + 2, [(1.0, 1, 0, 0)], 1, p1)
                    _, a2 = orc.elvec_h1(dim, fo, vn, pm, ls, dt, 2 * fo + 2, [(1.0, 1, 0, 0), (2.0, 1, 1, 0)])
                    assert np.abs(a1 - a2).max() <= 1e-13 * max(np.abs(a2).max(), 1e-3)
    assert nchk > 60


@pytest.mark.parametrize("dim", [2, 3])
def test_elmat_with_gridfunction_coefficient_pins(dim):
    """orc_elmat_h1_cf against pinned pieces: (a) constant coefficient c -> c * orc_elmat_h1; (b) mass form: the vertex functions are a
    partition of unity, so the row sums over the vertex columns are the element vector of the same coefficient (orc_elvec_h1_cf);
    (c) stiffness form applied to the linear function x_d: K_cf x_d = int gf d_d phi_i = the mass-form trick on the derivative side,
    checked through symmetry and K_cf * 1 = 0"""
    rng = np.random.default_rng(77 + dim)
    nchk = 0
    for trial in range(30):
        pm = rng.uniform(-1, 1, size=(dim + 1, dim))
        if abs(np.linalg.det((pm[:dim] - pm[dim]).T)) < 0.05:
            continue
        ls = rng.uniform(-1, 1, size=dim + 1) + (0.0 if trial % 3 else 2.0)
        vn = np.arange(dim + 1, dtype=np.int32)
        for fo in (1, 2):
            for cfo in (1, 2):
                ncf = orc.load().orc_h1_ndof(dim, cfo)
                cl = rng.uniform(-1, 1, ncf)
                for dt in (0, 1, 2):
                    n0, base = orc.elmat_h1(dim, fo, vn, pm, ls, dt, 2 * fo, 1.3, 0.7)
                    const = np.zeros(ncf); const[:dim + 1] = 2.5                  # hierarchical basis: the constant 2.5
                    n1, em_c = orc.elmat_h1_cf(dim, fo, vn, pm, ls, dt, 2 * fo, 1.3, 0.7, cfo, const)
                    assert max(n0, 0) == max(n1, 0)
                    if n0 <= 0:
                        continue
                    nchk += 1
                    assert np.abs(em_c - 2.5 * base).max() <= 1e-13 * np.abs(base).max()
                    _, mass = orc.elmat_h1_cf(dim, fo, vn, pm, ls, dt, 2 * fo, 1.0, 0.0, cfo, cl)
                    _, ev = orc.elvec_h1_cf(dim, fo, vn, pm, ls, dt, 2 * fo, [(1.0, 0, 0, 0)], cfo, cl)
                    assert np.abs(mass[:, :dim + 1].sum(axis=1) - ev).max() <= 1e-13 * max(np.abs(ev).max(), 1e-3)
                    _, stiff = orc.elmat_h1_cf(dim, fo, vn, pm, ls, dt, 2 * fo, 0.0, 1.0, cfo, cl)
                    assert np.abs(stiff - stiff.T).max() <= 1e-13 * np.abs(stiff).max()
                    assert np.abs(stiff[:, :dim + 1].sum(axis=1)).max() <= 1e-12 * np.abs(stiff).max()
    assert nchk > 100
