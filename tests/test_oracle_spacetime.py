"""Space-time cut rule: the C restatement (oracle/ngsx_oracle.c:orc_st_cut_rule) against
  (a) the reference's own spacetimecutrule.cpp + NodalTimeFE compiled in place (oracle/_ref), on randomised elements,
  (b) the reference's known-answer tests tests/pytests/test_spacetimecutrule.py:12-68 restated without NGSolve."""
import numpy as np
import pytest

from oracle import orc, refload

NEG, POS, IF = 0, 1, 2
needs_ref = pytest.mark.skipif(not refload.available(), reason="compiled reference (oracle/_ref) not built")


def _rand_element(rng, dim):
    pm = rng.uniform(-1, 1, size=(dim + 1, dim))
    while abs(np.linalg.det((pm[:dim] - pm[dim]).T)) < 0.05:
        pm = rng.uniform(-1, 1, size=(dim + 1, dim))
    return pm


@needs_ref
@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("ntd", [1, 2, 3])
def test_st_rule_vs_compiled_reference(dim, ntd):
    rng = np.random.default_rng(20241104 + 10 * dim + ntd)
    ncut = 0
    for trial in range(120):
        pm = _rand_element(rng, dim)
        ls = rng.uniform(-1, 1, size=(ntd, dim + 1)) + rng.uniform(-0.6, 0.6)
        if trial % 7 == 0 and ntd == 3:
            ls[1] = 0.5 * (ls[0] + ls[2])                 # linear in time on an order-2 time element: the reference's a == 0 branch
        for dt in (NEG, POS, IF):
            for ot, os_ in ((0, 0), (2, 2), (4, 3)):
                n1, p1, t1, w1 = orc.st_cut_rule(dim, pm, ls, dt, ot, os_)
                n2, p2, t2, w2 = orc.st_cut_rule(dim, pm, ls, dt, ot, os_, impl="ref")
                assert max(n1, 0) == max(n2, 0), (trial, dt, ot, os_)
                if n1 > 0:
                    ncut += 1
                    assert np.array_equal(t1, t2) or np.abs(t1 - t2).max() < 1e-14
                    assert np.abs(p1 - p2).max() < 1e-12
                    assert np.abs(w1 - w2).max() <= 1e-12 * max(np.abs(w2).max(), 1e-300)
    assert ncut > 200


def _unit_square():
    vert = np.array([[0, 0], [1, 0], [0, 1], [1, 1]], dtype=float)
    el2v = np.array([[0, 1, 2], [3, 2, 1]], dtype=np.int32)                  # NGS/python/meshes.py:109-117 (nx = ny = 1)
    return vert, el2v


@pytest.mark.parametrize("mono,ref,os_,ot", [([[1, 0, 0, 0, 1]], 0.5, 0, 1),
                                             ([[1, 0, 0, 0, 3]], 0.25, 0, 3),
                                             ([[1, 0, 0, 0, 0], [-3, 0, 0, 0, 1], [3, 0, 0, 0, 2], [-1, 0, 0, 0, 3]], 0.25, 0, 3),
                                             ([[1, 1, 0, 0, 0]], 0.5, 1, 0),
                                             ([[1, 2, 0, 0, 2], [1, 0, 2, 0, 2]], 2 / 9, 2, 2)])
def test_spacetime_integrate_no_cut(mono, ref, os_, ot):
    """tests/pytests/test_spacetimecutrule.py:12-36 (triangles): tref, tref^3, (1-tref)^3, x, tref^2 (x^2+y^2) on the unit square"""
    vert, el2v = _unit_square()
    ls = -np.ones((2, 4))
    val, _ = orc.st_integrate_mesh(2, el2v, vert, ls, NEG, os_, ot, mono)
    assert abs(val - ref) < 5e-15


@pytest.mark.parametrize("dt,ref", [(POS, 1 / 8), (NEG, 7 / 8), (IF, 1 / 2)])
def test_spacetime_straight_cut_2Dplus1D(dt, ref):
    """tests/pytests/test_spacetimecutrule.py:39-68: level set 1 - 2x - 2t, order = time_order = 0"""
    vert, el2v = _unit_square()
    ls = np.array([1 - 2 * vert[:, 0], 1 - 2 * vert[:, 0] - 2])
    val, _ = orc.st_integrate_mesh(2, el2v, vert, ls, dt, 0, 0)
    assert abs(val - ref) < 5e-15


@needs_ref
@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("ntd", [4, 5, 6])
def test_st_rule_bisection_orders_vs_compiled_reference(dim, ntd):
    """time elements of order 3..5: Gauss-Lobatto nodes (spacetime/SpaceTimeFE.cpp:150-173) and the subdivision + secant / bisection
    root search (cutint/spacetimecutrule.cpp:46-95) against the compiled reference"""
    rng = np.random.default_rng(777 + 10 * dim + ntd)
    ncut = 0
    for trial in range(60):
        pm = _rand_element(rng, dim)
        ls = rng.uniform(-1, 1, size=(ntd, dim + 1)) + rng.uniform(-0.4, 0.4)
        for dt in (NEG, POS, IF):
            for ot, os_ in ((0, 0), (3, 2)):
                n1, p1, t1, w1 = orc.st_cut_rule(dim, pm, ls, dt, ot, os_)
                n2, p2, t2, w2 = orc.st_cut_rule(dim, pm, ls, dt, ot, os_, impl="ref")
                assert max(n1, 0) == max(n2, 0), (trial, dt, ot, os_)
                if n1 > 0:
                    ncut += 1
                    assert np.abs(t1 - t2).max() < 1e-13
                    assert np.abs(p1 - p2).max() < 1e-11
                    assert np.abs(w1 - w2).max() <= 1e-11 * max(np.abs(w2).max(), 1e-300)
    assert ncut > 100


@needs_ref
@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("fo,ntd_fe,ntd", [(1, 1, 2), (1, 2, 2), (2, 2, 3), (1, 3, 3), (2, 3, 2), (1, 4, 4)])
def test_st_elmat_vs_compiled_spacetimefe(dim, fo, ntd_fe, ntd):
    """space-time element matrices: the restatement (orc_st_elmat_h1) against the reference's own SpaceTimeFE (CalcShape / CalcDShape /
    CalcDtShape, spacetime/SpaceTimeFE.cpp:29-113) over H1HighOrderFE x NodalTimeFE on the points of its SpaceTimeCutIntegrationRule"""
    rng = np.random.default_rng(4242 + 100 * dim + 10 * fo + ntd_fe)
    nchk = 0
    for trial in range(25):
        pm = _rand_element(rng, dim)
        ls = rng.uniform(-1, 1, size=(ntd, dim + 1)) + rng.uniform(-0.5, 0.5)
        for dt in (NEG, POS, IF):
            n1, m1 = orc.st_elmat(dim, pm, ls, fo, ntd_fe, dt, 2 * (ntd_fe - 1) + 1, 2 * fo, 1.3, 0.7, -0.9)
            n2, m2 = orc.st_elmat(dim, pm, ls, fo, ntd_fe, dt, 2 * (ntd_fe - 1) + 1, 2 * fo, 1.3, 0.7, -0.9, impl="ref")
            assert max(n1, 0) == max(n2, 0)
            if n1 > 0:
                nchk += 1
                assert np.abs(m1 - m2).max() <= 1e-12 * np.abs(m2).max()
                if ntd_fe > 1:
                    assert np.abs(m1 - m1.T).max() > 1e-6 * np.abs(m1).max()       # the dt(u) v term is there
    assert nchk > 20


def test_st_elmat_closed_forms():
    """uncut element, level set constant in time: the block is (time mass) x (spatial mass / stiffness) + (time advection) x (spatial mass)"""
    pm = np.array([[0.0, 0.0], [1.0, 0.0], [0.0, 1.0]])
    ls = -np.ones((2, 3))
    _, m = orc.st_elmat(2, pm, ls, 1, 2, NEG, 2, 2, 1.0, 0.0, 0.0)
    ms = np.array([[2, 1, 1], [1, 2, 1], [1, 1, 2]]) / 24.0                      # P1 mass matrix of the reference triangle
    mt = np.array([[2, 1], [1, 2]]) / 6.0                                          # P1 mass matrix of [0, 1]
    assert np.abs(m - np.kron(mt, ms)).max() < 1e-15
    _, a = orc.st_elmat(2, pm, ls, 1, 2, NEG, 2, 2, 0.0, 0.0, 1.0)
    ct = np.array([[-1, 1], [-1, 1]]) / 2.0                                        # int tau_j tau_j'' on [0, 1]
    assert np.abs(a - np.kron(ct, ms)).max() < 1e-15


@needs_ref
def test_st_cutinfo_element_vs_compiled_reference():
    """space-time CutInformation (xfem/cutinfo.cpp:79-111 with time_order >= 0): part volumes from the compiled reference's rules"""
    rng = np.random.default_rng(99)
    seen = set()
    for trial in range(200):
        dim = 2 + trial % 2
        pm = _rand_element(rng, dim)
        ntd = 2 + trial % 3
        ls = rng.uniform(-1, 1, size=(ntd, dim + 1)) + rng.uniform(-1.2, 1.2)
        d, r = orc.st_cutinfo_element(dim, pm, ls, 2)
        pv = [0.0, 0.0]
        for np_ in (POS, NEG):
            n, _, _, w = orc.st_cut_rule(dim, pm, ls, np_, 2, 0, impl="ref")
            pv[np_] = w.sum() if n > 0 else 0.0
        want = (IF if pv[POS] > 0 else NEG) if pv[NEG] > 0 else POS
        assert d == want
        assert abs(r - pv[NEG] / (pv[NEG] + pv[POS])) < 1e-13
        seen.add(d)
    assert seen == {NEG, POS, IF}
