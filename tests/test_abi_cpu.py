"""CPU-side checks of the product library: it loads, exports every symbol include/ngsx_abi.h declares, its host-only
entry points work (greedy group colouring), argument errors come back as statuses, and the host-side mirror of the
reference's Python surface parses the tutorial forms.  No kernel is launched here."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from ngsxfem_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _lib.load()


def test_exports_every_declared_symbol(lib):
    from ngsxfem_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "ngsx_abi.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ngsx_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/ngsx_abi.h but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert lib.ngsx_abi_version() == 1


def test_no_oracle_in_product():
    """the product must never import / link the oracle"""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "ngsxfem_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "ngsx_oracle" not in txt, f


def test_struct_layouts(lib):
    from ngsxfem_b200 import _lib
    assert ctypes.sizeof(_lib.Mesh) == 40
    assert ctypes.sizeof(_lib.FormH1) == 48
    assert ctypes.sizeof(_lib.XTerm) == 32
    assert ctypes.sizeof(_lib.Poly) == 8 + 8 * 8 + 8 * 3 * 4


def test_argument_errors_are_statuses(lib):
    from ngsxfem_b200 import _lib
    m = _lib.Mesh(5, 0, 1, 1, None, None)
    assert lib.ngsx_classify(ctypes.byref(m), None, None, None, None, None, None) == -1
    assert b"dim" in lib.ngsx_last_error()
    with pytest.raises(_lib.NgsxError):
        _lib.check(lib.ngsx_slot_map(-1, 3, None, None, None, None, None))
    with pytest.raises(_lib.NgsxError):
        _lib.make_poly([(1.0, 0.5, 0, 0)])


@pytest.mark.parametrize("group", [1, 4, 6])
def test_color_groups_host(lib, group):
    from ngsxfem_b200 import meshes
    m = meshes.MakeStructured3DMesh(5)
    e2d, ndof = m.h1_dofs(2)
    tab = e2d.contiguous()
    n, nd = tab.shape
    perm = torch.empty(n, dtype=torch.int64)
    cptr = torch.zeros(257, dtype=torch.int64)
    nc = ctypes.c_int32(0)
    assert lib.ngsx_color_groups(n, nd, tab.data_ptr(), ndof, group, perm.data_ptr(), cptr.data_ptr(), 256, ctypes.byref(nc)) == 0
    nc = nc.value
    cp = cptr[: nc + 1].numpy()
    assert cp[0] == 0 and cp[-1] == n and (np.diff(cp) > 0).all()
    assert np.array_equal(np.sort(perm.numpy()), np.arange(n))                     # a permutation
    t = tab.numpy()[perm.numpy()]
    for c in range(nc):
        seg = t[cp[c]:cp[c + 1]]
        ng = (len(seg) + group - 1) // group
        owner = {}
        for g in range(ng):                                                        # the kernel's windows inside a colour
            for d in np.unique(seg[g * group:(g + 1) * group]):
                assert owner.setdefault(d, g) == g, "two groups of one colour share a dof"
    if group == 1:
        assert nc >= 24          # a Kuhn-mesh vertex is shared by up to 24 tets
    # too small capacity is reported, not overrun
    nc2 = ctypes.c_int32(0)
    assert lib.ngsx_color_groups(n, nd, tab.data_ptr(), ndof, group, perm.data_ptr(), cptr.data_ptr(), 2, ctypes.byref(nc2)) == -5


def test_host_mirror_parses_tutorial_forms():
    from ngsxfem_b200 import _lib, xfem
    from ngsxfem_b200.ngs import Expr, LevelsetNormal, grad, specialcf, x, y
    u, v = Expr.atom("u"), Expr.atom("v")
    h, n = specialcf.mesh_size, LevelsetNormal()
    b = xfem.SymbolicCutBFI(None, domain_type=xfem.NEG, form=grad(u) * grad(v))
    assert (b.form.c_stiff, b.form.c_mass, b.intorder(2), b.intorder(3)) == (1.0, 0.0, 2, 4)    # symboliccutbfi.cpp:97-107
    b = xfem.SymbolicCutBFI(None, domain_type=xfem.NEG, form=grad(u) * grad(v) + u * v)
    assert b.intorder(2) == 4
    b = xfem.SymbolicCutBFI(None, domain_type=xfem.NEG, form=u * v, force_intorder=7)
    assert b.intorder(2) == 7
    nit = -grad(u) * n * v - grad(v) * n * u + (90 / h) * u * v                                 # fictdom_ghostpen.py:88-93
    b = xfem.SymbolicCutBFI(None, domain_type=xfem.IF, form=nit)
    assert (b.form.c_nitsche, b.form.c_mass, b.form.hpow_mass) == (-1.0, 90.0, -1)
    gp = xfem.SymbolicFacetPatchBFI(form=0.1 * 1.0 / h * 1.0 / h * (u - u.Other()) * (v - v.Other()), skeleton=False)
    assert (gp.coef, gp.hpow) == (0.1, -2)
    lf = xfem.SymbolicCutLFI(None, form=(1 + 2 * x * y) * v)
    assert lf.coef.terms == {(0, 0, 0): 1.0, (1, 1, 0): 2.0} and lf.intorder(2) == 4
    for bad in (lambda: xfem.SymbolicCutBFI(None, form=grad(u) * n * v),                        # unsymmetric Nitsche: unsupported -> error
                lambda: xfem.SymbolicCutBFI(None, form=u * v, element_boundary=True),
                lambda: xfem.SymbolicCutBFI(None, form=u * v, time_order=2),
                lambda: xfem.SymbolicFacetPatchBFI(form=u * v, skeleton=False),
                lambda: xfem.SymbolicFacetPatchBFI(form=(u - u.Other()) * (v - v.Other()), skeleton=True),
                lambda: xfem.SymbolicCutLFI(None, form=u * v)):
        with pytest.raises(_lib.NgsxError):
            bad()
    assert (xfem.NEG, xfem.POS, xfem.IF) == (0, 1, 2) and int(xfem.HASNEG) == 5 and int(xfem.ANY) == 7   # utils/ngsxstd.hpp


def test_xfem_form_expansion():
    """f-2 host logic: the Nitsche-XFEM forms of py_tutorials/nxfem.py:77-107 expand into the product terms of ngsx_xterm with
    the reference's differential orders (xfem/symboliccutbfi.cpp:97-101, xfem/xfemdiffops.hpp:24)"""
    from ngsxfem_b200 import _lib, ngs, xfes
    from ngsxfem_b200.ngs import Expr
    u_std, v_std = Expr.atom("u"), Expr.atom("v")
    u_x, v_x = xfes.XProxy("u"), xfes.XProxy("v")
    alpha = [1.0, 2.0]
    n, h = ngs.LevelsetNormal(), ngs.specialcf.mesh_size
    k0 = xfes.CutRatioGF(None)
    kappa = (k0, 1.0 - k0)
    u = [u_std + op(u_x) for op in (xfes.neg, xfes.pos)]
    v = [v_std + op(v_x) for op in (xfes.neg, xfes.pos)]
    gu = [ngs.grad(u_std) + op(u_x) for op in (xfes.neg_grad, xfes.pos_grad)]
    gv = [ngs.grad(v_std) + op(v_x) for op in (xfes.neg_grad, xfes.pos_grad)]
    terms, do = xfes.xterms_of(alpha[1] * gu[1] * gv[1])
    assert do == (1, 1) and sorted(terms) == sorted([(2.0, 0, 0, 4, bu, bv) for bu in (0, 2) for bv in (0, 2)])
    fu = sum([-kappa[i] * alpha[i] * gu[i] * n for i in (0, 1)])
    fv = sum([-kappa[i] * alpha[i] * gv[i] * n for i in (0, 1)])
    stab = 20 * (alpha[0] + alpha[1]) / h
    terms, do = xfes.xterms_of(fu * (v[0] - v[1]) + fv * (u[0] - u[1]) + stab * (u[0] - u[1]) * (v[0] - v[1]))
    assert do == (0, 0) and len(terms) == 24
    # the jump [v] = neg(v_x) - pos(v_x): no term may test with the dofs of Vh, and the penalty carries h^-1
    assert all(bv != 0 for c, p, k, pair, bu, bv in terms if pair in (0, 1))
    assert sorted((c, p) for c, p, k, pair, bu, bv in terms if pair == 0) == [(-60.0, -1), (-60.0, -1), (60.0, -1), (60.0, -1)]
    assert xfes.has_x_atoms(fu * v[0]) and not xfes.has_x_atoms(ngs.grad(u_std) * ngs.grad(v_std))
    with pytest.raises(_lib.NgsxError):
        xfes.xterms_of(k0 * k0 * u[0] * v[0])
    with pytest.raises(_lib.NgsxError):
        xfes.neg(u_std)


def test_xfem_linearform_parse():
    """f-2 host logic: f * (v_std + neg(v_x)) (py_tutorials/nxfem.py:109-110) -> one scalar per test block and the polynomial f;
    intorder = 2 order (xfem/symboliccutlfi.cpp:85); forms with a trial function or a power of h are refused"""
    import pytest
    from ngsxfem_b200 import ngs, xfem, xfes
    from ngsxfem_b200._lib import NgsxError
    from ngsxfem_b200.ngs import Expr
    v_std, v_x = Expr.atom("v"), xfes.XProxy("v")
    f0 = 1.5 + ngs.x * ngs.y
    li = xfem.SymbolicLFI(levelset_domain={"levelset": None, "domain_type": xfem.NEG}, form=f0 * (v_std + xfes.neg(v_x)))
    assert li.blocks == [1.0, 1.0, 0.0] and li.has_x and li.intorder(2) == 4
    assert li.poly.terms == {(0, 0, 0): 1.5, (1, 1, 0): 1.0}
    lp = xfem.SymbolicLFI(levelset_domain={"levelset": None, "domain_type": xfem.POS}, form=2.0 * (v_std + xfes.pos(v_x)))
    assert lp.blocks == [2.0, 0.0, 2.0]
    plain = xfem.SymbolicLFI(levelset_domain={"levelset": None, "domain_type": xfem.NEG}, form=f0 * v_std)
    assert plain.blocks == [1.0, 0.0, 0.0] and not plain.has_x
    with pytest.raises(NgsxError):
        xfem.SymbolicLFI(levelset_domain={"levelset": None, "domain_type": xfem.NEG}, form=Expr.atom("u") * v_std)
    with pytest.raises(NgsxError):
        xfem.SymbolicLFI(levelset_domain={"levelset": None, "domain_type": xfem.NEG}, form=ngs.specialcf.mesh_size * v_std)


def test_dcut_symbols_build_the_same_integrators():
    """form * dCut(lset, NEG, order=...) and form * dx(levelset_domain=...) are spellings of SymbolicBFI / SymbolicLFI(levelset_domain=...)"""
    from ngsxfem_b200 import ngs, xfem
    from ngsxfem_b200.ngs import Expr
    u, v = Expr.atom("u"), Expr.atom("v")
    form = ngs.grad_of(u) * ngs.grad_of(v) + u * v if hasattr(ngs, "grad_of") else u * v
    integ = form * xfem.dCut(None, xfem.NEG, order=3)
    assert isinstance(integ, xfem.CutIntegrand) and len(integ.parts) == 1
    f, sym = integ.parts[0]
    assert f is form and sym.levelset_domain == {"levelset": None, "domain_type": xfem.NEG, "force_intorder": 3}
    two = form * xfem.dCut(None, xfem.NEG) + 2.0 * form * xfem.dx(levelset_domain={"levelset": None, "domain_type": xfem.POS})
    assert [p[1].levelset_domain["domain_type"] for p in two.parts] == [xfem.NEG, xfem.POS]
    rhs = (1.0 + ngs.x) * v * xfem.dCut(None, xfem.IF)
    assert rhs.parts[0][0].poly.terms == {(0, 0, 0): 1.0, (1, 0, 0): 1.0}


def test_deformation_host_logic():
    """row f-4 host side (no GPU): mesh.SetDeformation validates the space (vector H1 of order 1..2 with dim = mesh.dim), the
    ngsx_deformation struct has the C layout of include/ngsx_abi.h, integrators and differential symbols carry their deformation,
    element blocks inherit the mesh's, and nothing reaches a kernel with host tensors"""
    import ctypes
    import pytest
    from ngsxfem_b200 import _lib, meshes, ngs, xfem
    from ngsxfem_b200.ngs import Expr
    assert ctypes.sizeof(_lib.Deformation) == 32 and _lib.Deformation.el2dof.offset == 16 and _lib.Deformation.coef.offset == 24
    mesh = ngs.Mesh(meshes.MakeStructured2DMesh(3), device="cpu")
    gf = ngs.GridFunction(ngs.H1(mesh, order=1))
    for bad in (gf, ngs.GridFunction(ngs.H1(mesh, order=3, dim=2)), ngs.GridFunction(ngs.H1(mesh, order=2, dim=3))):
        with pytest.raises(_lib.NgsxError):
            mesh.SetDeformation(bad)
    defo = ngs.GridFunction(ngs.H1(mesh, order=2, dim=2))
    assert tuple(defo.vec.shape) == (defo.space.ndof, 2)
    mesh.SetDeformation(defo)
    blk = mesh.element_block(2, 7)
    assert blk.deformation is defo and blk._block == (2, 7)
    with pytest.raises(_lib.NgsxError):                                    # host tensors never reach the library
        mesh.c_deformation()
    mesh.UnsetDeformation()
    assert mesh.deformation is None and mesh.c_deformation() is None
    u, v = Expr.atom("u"), Expr.atom("v")
    bfi = xfem.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xfem.NEG}, form=u * v, deformation=defo)
    lfi = xfem.SymbolicLFI(levelset_domain={"levelset": gf, "domain_type": xfem.NEG}, form=1.0 * v, deformation=defo)
    assert bfi.deformation is defo and lfi.deformation is defo
    sym = xfem.dCut(gf, xfem.IF, deformation=defo)
    assert sym.deformation is defo and sym(order=3).deformation is defo and xfem.dx(levelset_domain=sym.levelset_domain).deformation is None
    a = xfem.BilinearForm(ngs.H1(mesh, order=2))
    a += (u * v) * sym
    assert a.integrators[0].deformation is defo and a._deformation_of(a.integrators[0]) is defo


def test_weighted_element_blocks():
    """sweep blocks of the multi-GPU step: contiguous, tiling, balanced by weight; degenerate inputs"""
    import torch
    from ngsxfem_b200 import partition
    w = torch.ones(1000, dtype=torch.float64)
    w[400:600] += 40.0
    for world in (1, 2, 3, 8):
        b = partition.element_blocks_weighted(w, world)
        assert len(b) == world and b[0][0] == 0 and b[-1][1] == 1000 and all(b[r][1] == b[r + 1][0] for r in range(world - 1))
        shares = [float(w[lo:hi].sum() / w.sum()) for lo, hi in b]
        assert max(shares) <= 1.0 / world + 0.01
    assert partition.element_blocks_weighted(torch.ones(3, dtype=torch.float64), 8)[-1][1] == 3
    assert partition.element_blocks_weighted(torch.zeros(0, dtype=torch.float64), 2) == [(0, 0), (0, 0)]
