"""Host-side mirror of the reference's Python surface for the hot path (same names, argument meaning and error
behaviour): python/__init__.py:107-378, xfem/python_xfem.cpp:50-815, utils/python_utils.cpp:21-153,
cutint/python_cutint.cpp:20-81.  Every numerical step is one batched call into the C ABI (include/ngsx_abi.h);
this file only validates arguments, owns device buffers (torch tensors) and caches the symbolic phase
(sparsity graph, colouring, slot map).  Anything the CUDA path does not cover raises -- there is no CPU fallback."""
import ctypes
from typing import Dict, List, Optional

import torch

from . import _lib
from ._lib import NEG, POS, IF, NgsxError, check
from .ngs import (BND, VOL, AnalyticLevelset, BitArray, CSRMatrix, Expr, GridFunction, H1, Mesh, Poly, _ptr, _stream,
                  require_cuda)

# enums exported by python/python_ngsxfem.cpp:14-37
NO, CDOM_NEG, CDOM_POS, UNCUT, CDOM_IF, HASNEG, HASPOS, ANY = range(8)
FIRST, OPTIMAL, FALLBACK = 0, 1, 2      # QUAD_DIRECTION_POLICY (quads/hexes only; ignored on simplices)
_DT2CDT = {NEG: CDOM_NEG, POS: CDOM_POS, IF: CDOM_IF}


def _lib_():
    return _lib.load()


def _lset_vec(levelset) -> torch.Tensor:
    """CF2GFForStraightCutRule (cutint/xintegration.cpp:91-103): only a GridFunction of order <= 1 on an H1 space takes the
    straight-cut path.  The legacy subdivision path for general CoefficientFunctions is out of scope -> error."""
    if isinstance(levelset, GridFunction) and levelset.space.order == 1:
        return levelset.vec
    raise NgsxError(-2, "level set must be a P1 GridFunction (InterpolateToP1 first); the legacy subdivlvl path of "
                        "cutint/xintegration.cpp:696-1007 is outside the CUDA path")


# ------------------------------------------------------------------------------------------------------------------
def InterpolateToP1(coef, gf_p1: GridFunction, eps_perturbation: float = 1e-14, heapsize: int = 1000000):
    """utils/python_utils.cpp:21-81 -> InterpolateP1::Do (utils/p1interpol.cpp:25-96)."""
    mesh = gf_p1.space.mesh
    out = gf_p1.vec
    require_cuda(out, "gf_p1")
    L = _lib_()
    if isinstance(coef, GridFunction):                               # GridFunction branch: copy vertex dofs, clamp
        src = coef.vec[: mesh.nv].contiguous()
        check(L.ngsx_interpolate_p1_clamp(mesh.nv, _ptr(src), eps_perturbation, _ptr(out), _stream()))
    elif isinstance(coef, AnalyticLevelset):
        par = (ctypes.c_double * 8)(*coef.params)
        check(L.ngsx_interpolate_p1_cf(ctypes.byref(mesh.c_struct()), coef.kind, par, eps_perturbation, _ptr(out), _stream()))
    elif isinstance(coef, Poly) and coef.is_affine():
        t = coef.terms
        par = (ctypes.c_double * 8)(t.get((0, 0, 0), 0.0), t.get((1, 0, 0), 0.0), t.get((0, 1, 0), 0.0), t.get((0, 0, 1), 0.0), 0, 0, 0, 0)
        check(L.ngsx_interpolate_p1_cf(ctypes.byref(mesh.c_struct()), _lib.LSET_AFFINE, par, eps_perturbation, _ptr(out), _stream()))
    elif callable(coef):                                             # arbitrary CF: evaluated at the vertices by the host (plumbing)
        cols = [mesh.vert[:, d] for d in range(mesh.dim)]
        vals = coef(*cols).to(torch.float64).contiguous()
        check(L.ngsx_interpolate_p1_clamp(mesh.nv, _ptr(vals), eps_perturbation, _ptr(out), _stream()))
    else:
        raise NgsxError(-2, "InterpolateToP1: unsupported coefficient")


# ------------------------------------------------------------------------------------------------------------------
class CutInfo:
    """xfem/python_xfem.cpp:50-165 / CutInformation (xfem/cutinfo.cpp:11-188)."""

    def __init__(self, mesh: Mesh, levelset=None, time_order: int = -1, heapsize: int = 1000000):
        self.mesh = mesh
        dev = mesh.device
        self.nwords = (mesh.ne + 31) // 32
        self.words = torch.zeros((8, self.nwords), dtype=torch.int32, device=dev)
        self.dom = torch.zeros(mesh.ne, dtype=torch.int8, device=dev)
        self.ratio = torch.zeros(mesh.ne, dtype=torch.float64, device=dev)
        self.err = torch.zeros(2, dtype=torch.int64, device=dev)
        self.cut_neighboring_node: Dict[str, BitArray] = {}
        self.dom_of_node: Dict[str, torch.Tensor] = {}
        if levelset is not None:
            self.Update(levelset, time_order, heapsize)

    def Mesh(self):
        return self.mesh

    def Update(self, levelset, time_order: int = -1, heapsize: int = 1000000, node_marks: bool = False):
        L = _lib_()
        if time_order >= 0:                                          # space-time prisms: part volumes from the space-time rules
            from . import spacetime
            self.dom, self.ratio = spacetime.classify(levelset, self.mesh, time_order)
            d = self.dom
            for cdt, mask in ((1, d == NEG), (2, d == POS), (3, (d == NEG) | (d == POS)), (4, d == IF), (5, (d == NEG) | (d == IF)),
                              (6, (d == POS) | (d == IF)), (7, torch.ones_like(d, dtype=torch.bool))):
                self.words[cdt] = BitArray.from_bool(mask).words
            self._bnd = None
            return self
        lv = _lset_vec(levelset)
        if getattr(self.mesh, "tensor", False):                      # quadrilaterals / hexahedra (row f-5)
            check(L.ngsx_q_classify(ctypes.byref(self.mesh.c_struct()), _ptr(lv), _ptr(self.dom), _ptr(self.ratio), _ptr(self.err), _stream()))
            d = self.dom
            for cdt, mask in ((1, d == NEG), (2, d == POS), (3, (d == NEG) | (d == POS)), (4, d == IF), (5, (d == NEG) | (d == IF)),
                              (6, (d == POS) | (d == IF)), (7, torch.ones_like(d, dtype=torch.bool))):
                self.words[cdt] = BitArray.from_bool(mask).words
            return self
        check(L.ngsx_classify(ctypes.byref(self.mesh.c_struct()), _ptr(lv), _ptr(self.words), _ptr(self.dom), _ptr(self.ratio),
                              _ptr(self.err), _stream()))
        self._bnd = None                                             # the BND pass (xfem/cutinfo.cpp:47, 67) runs when it is asked for
        self._bnd_lv = lv
        if node_marks:
            self._node_marks()
        return self

    def check_degenerate(self):
        """the reference throws from inside Update on these level sets; here the kernel counts them"""
        n, first = self.err.tolist()
        if n:
            raise NgsxError(-4, f"{n} element(s) carry a level set the reference rejects (first: element {first}): "
                                "'You tried to cut a simplex with a plain geometry lset function'")

    def _node_marks(self):
        L = _lib_()
        m = self.mesh
        tables = {"vertex": (m.el2v, m.nv)}
        e2v, el2e = m.edges()
        tables["edge"] = (el2e.to(torch.int32).contiguous(), e2v.shape[0])
        if m.dim == 3:
            f2v, el2f = m.faces()
            tables["face"] = (el2f.to(torch.int32).contiguous(), f2v.shape[0])
        for name, (tab, nn) in tables.items():
            ba = BitArray(nn, device=m.device)
            don = torch.full((nn,), IF, dtype=torch.int32, device=m.device)
            check(L.ngsx_node_marks(m.ne, tab.shape[1], _ptr(tab), _ptr(self.dom), _ptr(self.ratio), nn, 0, _ptr(ba.words), _ptr(don), _stream()))
            self.cut_neighboring_node[name], self.dom_of_node[name] = ba, don
        ba = BitArray(m.ne, device=m.device)
        don = torch.full((m.ne,), IF, dtype=torch.int32, device=m.device)
        check(L.ngsx_node_marks(m.ne, 0, None, _ptr(self.dom), _ptr(self.ratio), m.ne, 1, _ptr(ba.words), _ptr(don), _stream()))
        self.cut_neighboring_node["element"], self.dom_of_node["element"] = ba, don

    def GetElementsOfType(self, domain_type=IF, VOL_or_BND=VOL) -> BitArray:
        """domain_type: DOMAIN_TYPE (NEG/POS/IF) or COMBINED_DOMAIN_TYPE (HASNEG, ...; the module constants are wrapped ints so
        the two reference enums, which share small integers, stay distinguishable) -- xfem/python_xfem.cpp:91-118."""
        cdt = domain_type.cdt if isinstance(domain_type, _CDT) else _DT2CDT[int(domain_type)]
        if VOL_or_BND != VOL:
            words, _, _ = self._boundary()
            return BitArray(self.mesh.boundary_elements().shape[0], words[cdt])
        return BitArray(self.mesh.ne, self.words[cdt])

    def _boundary(self):
        """CutInformation::Update's BND pass (xfem/cutinfo.cpp:47, 67-120): the same classification on the boundary elements"""
        if getattr(self, "_bnd", None) is None:
            if getattr(self.mesh, "tensor", False):
                raise NgsxError(-2, "boundary elements of quadrilateral / hexahedral meshes are outside the CUDA path")
            bel = self.mesh.boundary_elements()
            nbe, dev = bel.shape[0], self.mesh.device
            words = torch.zeros((8, (nbe + 31) // 32), dtype=torch.int32, device=dev)
            dom = torch.zeros(nbe, dtype=torch.int8, device=dev)
            ratio = torch.zeros(nbe, dtype=torch.float64, device=dev)
            err = torch.zeros(2, dtype=torch.int64, device=dev)
            check(_lib_().ngsx_classify_bnd(self.mesh.dim, nbe, _ptr(bel), _ptr(self._bnd_lv), _ptr(words), _ptr(dom), _ptr(ratio), _ptr(err), _stream()))
            self._bnd = (words, dom, ratio)
        return self._bnd

    def GetFacetsOfType(self, domain_type=IF) -> BitArray:
        # Appendix C.1: facets_of_domain_type is never updated by the reference: NEG all ones, POS/IF all zeros
        nf = self.mesh.nfacet
        ba = BitArray(nf, device=self.mesh.device)
        return ba.Set() if int(domain_type) == NEG and not isinstance(domain_type, _CDT) else ba

    def GetCutRatios(self, VOL_or_BND=VOL) -> torch.Tensor:
        return self.ratio if VOL_or_BND == VOL else self._boundary()[2]


def UpdateAndIntegrate(cutinfo: CutInfo, levelset, integrals, sync: bool = False):
    """CutInfo.Update(levelset) and Integrate(levelset_domain = {levelset, domain_type}, cf, order) for every (domain_type, order,
    cf) of `integrals` (at most 4) in ONE pass over the element table (ngsx_classify_integrate): the same results, bit for bit,
    as the separate calls.  Returns a device tensor with the integrals (a list of floats with sync = True)."""
    mesh = cutinfo.mesh
    if getattr(mesh, "tensor", False) or mesh.c_deformation() is not None:
        raise NgsxError(-2, "UpdateAndIntegrate: affine simplex meshes only (use the separate calls)")
    lv = _lset_vec(levelset)
    n = len(integrals)
    dts = (ctypes.c_int32 * max(n, 1))(*[int(t[0]) for t in integrals])
    orders = (ctypes.c_int32 * max(n, 1))(*[int(t[1]) for t in integrals])
    polys = (_lib.Poly * max(n, 1))(*[_poly_of(t[2] if len(t) > 2 else 1.0) for t in integrals])
    out = torch.zeros(max(n, 1), dtype=torch.float64, device=mesh.device)
    check(_lib_().ngsx_classify_integrate(ctypes.byref(mesh.c_struct()), _ptr(lv), _ptr(cutinfo.words), _ptr(cutinfo.dom), _ptr(cutinfo.ratio),
                                          _ptr(cutinfo.err), n, dts, orders, polys, _ptr(out), _stream()))
    cutinfo._bnd, cutinfo._bnd_lv = None, lv
    return out[:n].tolist() if sync else out[:n]


class _CDT(int):
    """COMBINED_DOMAIN_TYPE value"""
    def __new__(cls, v):
        o = int.__new__(cls, v)
        o.cdt = v
        return o


NO, CDOM_NEG, CDOM_POS, UNCUT, CDOM_IF, HASNEG, HASPOS, ANY = (_CDT(i) for i in range(8))


# ------------------------------------------------------------------------------------------------------------------
def GetFacetsWithNeighborTypes(mesh: Mesh, a: BitArray, bnd_val_a: bool = True, bnd_val_b: bool = True, use_and: bool = True,
                               b: Optional[BitArray] = None, heapsize: int = 1000000) -> BitArray:
    """xfem/python_xfem.cpp:168-214 / xfem/cutinfo.cpp:191-255"""
    if b is None:
        b = a
    _, f2e = mesh.facets()
    out = BitArray(f2e.shape[0], device=mesh.device)
    per = getattr(mesh, "periodic_facet", None)                      # MeshAccess::GetPeriodicFacet per facet (int32 [nf]) on periodic meshes
    if per is not None:
        check(_lib_().ngsx_facets_neighbor_types_periodic(f2e.shape[0], _ptr(f2e), _ptr(per), _ptr(a.words), _ptr(b.words), int(bnd_val_a),
                                                          int(bnd_val_b), int(use_and), _ptr(out.words), _stream()))
        return out
    check(_lib_().ngsx_facets_neighbor_types(f2e.shape[0], _ptr(f2e), _ptr(a.words), _ptr(b.words), int(bnd_val_a), int(bnd_val_b),
                                             int(use_and), _ptr(out.words), _stream()))
    return out


def GetElementsWithNeighborFacets(mesh: Mesh, a: BitArray, heapsize: int = 1000000) -> BitArray:
    """xfem/cutinfo.cpp:257-279"""
    _, f2e = mesh.facets()
    out = BitArray(mesh.ne, device=mesh.device)
    check(_lib_().ngsx_elements_with_neighbor_facets(f2e.shape[0], _ptr(f2e), _ptr(a.words), mesh.ne, _ptr(out.words), _stream()))
    return out


def GetDofsOfElements(space: H1, a: BitArray, heapsize: int = 1000000) -> BitArray:
    """xfem/cutinfo.cpp:281-304"""
    out = BitArray(space.ndof, device=space.mesh.device)
    check(_lib_().ngsx_dofs_of_elements(space.mesh.ne, space.nd, _ptr(space.el2dof), _ptr(a.words), space.ndof, _ptr(out.words), _stream()))
    return out


def GetDofsOfFacets(space: H1, a: BitArray, heapsize: int = 1000000) -> BitArray:
    """xfem/cutinfo.cpp:306-329: the dofs FESpace::GetDofNrs(NodeId(NT_FACET, f)) associates with the marked facets -- for an H1
    space the INNER dofs of the facet (edge dofs in 2D, face dofs in 3D; NGS/comp/h1hofespace.cpp GetDofNrs(NodeId)), not the
    vertex dofs.  Same kernel as GetDofsOfElements on the facet -> dof table."""
    mesh = space.mesh
    out = BitArray(space.ndof, device=mesh.device)
    nf = mesh.nfacet
    if mesh.dim == 2:
        per, base = space.order - 1, mesh.nv                         # edge f: dofs nv + f (order - 1) + k
    else:
        per = 1 if space.order >= 3 else 0                           # face f: one dof at order 3, after the vertex and edge dofs
        base = mesh.nv + mesh.edges()[0].shape[0] * (space.order - 1)
    if per <= 0 or nf == 0:
        return out
    f = torch.arange(nf, device=mesh.device, dtype=torch.int64)
    table = (base + f[:, None] * per + torch.arange(per, device=mesh.device)[None, :]).to(torch.int32).contiguous()
    check(_lib_().ngsx_dofs_of_elements(nf, per, _ptr(table), _ptr(a.words), space.ndof, _ptr(out.words), _stream()))
    return out


# ------------------------------------------------------------------------------------------------------------------
def _poly_of(cf):
    if cf is None:
        return _lib.make_poly(None)
    p = Poly.lift(cf)
    if p is None:
        raise NgsxError(-2, "only polynomial coefficient functions are on the CUDA path")
    return _lib.make_poly(p.as_terms())


def IntegrateX(lset, mesh: Mesh, cf=None, order: int = 5, domain_type=IF, subdivlvl: int = 0, time_order: int = -1,
               quad_dir_policy=OPTIMAL, heapsize: int = 1000000, sync: bool = True):
    """cutint/python_cutint.cpp:20-72.  Deterministic tree reduction on the device (the reference's AtomicAdd order is not)."""
    if subdivlvl != 0:
        raise NgsxError(-2, "subdivlvl > 0 (legacy integration) is outside the CUDA path")
    if time_order >= 0:                                              # space-time rule (cutint/xintegration.cpp:38-58)
        from . import spacetime
        return spacetime.integrate(lset, mesh, cf, order, domain_type, time_order, sync)
    lv = _lset_vec(lset)
    out = torch.zeros(1, dtype=torch.float64, device=mesh.device)
    from .ngs import _GfCoef
    if isinstance(cf, (GridFunction, _GfCoef)):                      # integrand = polynomial x H1 GridFunction (order 1-2), evaluated per point
        gfc, pol = (cf, 1.0) if isinstance(cf, GridFunction) else (cf.gf, cf.poly)
        cfs = gfc.space
        if getattr(mesh, "tensor", False) or mesh.c_deformation() is not None or not isinstance(cfs, H1) or getattr(cfs, "dim", 1) != 1 or \
                cfs.order not in (1, 2) or cfs.mesh is not mesh:
            raise NgsxError(-2, "a GridFunction integrand is served on affine simplex meshes for scalar H1 of order 1-2 (same mesh); no CPU fallback")
        poly = _poly_of(pol)
        check(_lib_().ngsx_integrate_cf(ctypes.byref(mesh.c_struct()), _ptr(lv), int(domain_type), order, ctypes.byref(poly), cfs.order,
                                        _ptr(cfs.el2dof), _ptr(gfc.vec), _ptr(out), _stream()))
        return float(out.item()) if sync else out
    poly = _poly_of(cf)
    if getattr(mesh, "tensor", False):                               # quadrilaterals / hexahedra (row f-5)
        if quad_dir_policy != OPTIMAL:
            raise NgsxError(-2, "quad_dir_policy FIRST / FALLBACK are outside the CUDA path (FIND_OPTIMAL, the reference's default, is)")
        err = torch.zeros(2, dtype=torch.int64, device=mesh.device)
        check(_lib_().ngsx_q_integrate(ctypes.byref(mesh.c_struct()), _ptr(lv), int(domain_type), order, ctypes.byref(poly), _ptr(out),
                                       _ptr(err), _stream()))
        if sync:
            n, first = err.tolist()
            if n:
                raise NgsxError(-4, f"{n} element(s) carry a level set the reference's quadrilateral rule rejects (first: element {first})")
        return float(out.item()) if sync else out
    defo = mesh.c_deformation()
    if defo is not None:                                             # mesh.SetDeformation(...): per-point geometry (row f-4)
        check(_lib_().ngsx_integrate_deformed(ctypes.byref(mesh.c_struct()), ctypes.byref(defo), _ptr(lv), int(domain_type), order,
                                              ctypes.byref(poly), _ptr(out), _stream()))
    else:
        check(_lib_().ngsx_integrate(ctypes.byref(mesh.c_struct()), _ptr(lv), int(domain_type), order, ctypes.byref(poly), _ptr(out), _stream()))
    return float(out.item()) if sync else out


def Integrate(levelset_domain=None, cf=None, mesh=None, order=5, time_order=-1, heapsize=1000000, **kwargs):
    """python/__init__.py:311-378: Integrate(levelset_domain={...}, cf=..., mesh=..., order=...); also Integrate(cf * dCut(...), mesh)"""
    if isinstance(levelset_domain, CutIntegrand):
        mesh = cf if (mesh is None and isinstance(cf, Mesh)) else mesh
        total = 0.0
        for f, sym in levelset_domain.parts:
            ld = sym.levelset_domain
            saved = mesh.deformation
            if sym.deformation is not None:                          # cf * dCut(..., deformation = gf): the symbol's deformation for this part
                mesh.SetDeformation(sym.deformation)
            try:
                total += Integrate(levelset_domain=ld, cf=f, mesh=mesh, order=ld.get("force_intorder", order), time_order=time_order)
            finally:
                mesh.deformation = saved
        return total
    if levelset_domain is None or not isinstance(levelset_domain, dict):
        raise NgsxError(-2, "only Integrate(levelset_domain=...) is provided (plain NGSolve Integrate is not part of the path)")
    if mesh is None:
        raise NgsxError(-1, "Integrate needs cf and mesh")
    ld = dict(levelset_domain)
    if "levelset" not in ld or "domain_type" not in ld:
        raise NgsxError(-1, "levelset_domain needs 'levelset' and 'domain_type'")            # python/__init__.py:281-307
    return IntegrateX(lset=ld["levelset"], mesh=mesh, cf=cf, order=order, domain_type=ld["domain_type"],
                      subdivlvl=ld.get("subdivlvl", 0), time_order=time_order,
                      quad_dir_policy=ld.get("quad_dir_policy", OPTIMAL), heapsize=heapsize)


# ------------------------------------------------------------------------------------------------------------------
# differential symbols: form * dCut(levelset, NEG) / form * dx(levelset_domain = {...}).  The reference at this revision only has
# the SymbolicBFI(levelset_domain = ...) spelling (python/__init__.py:107-279); later ngsxfem releases and the task statement
# name dCut / dx, so both spellings build the same integrators here.
# ------------------------------------------------------------------------------------------------------------------
class CutIntegrand:
    """sum of (form, differential symbol) pairs: what `a += ...`, `f += ...` and Integrate(...) accept"""

    def __init__(self, parts):
        self.parts = list(parts)

    def __add__(self, o):
        return CutIntegrand(self.parts + o.parts) if isinstance(o, CutIntegrand) else NotImplemented


class CutDifferentialSymbol:
    def __init__(self, levelset_domain, definedonelements=None, deformation=None):
        self.levelset_domain, self.definedonelements, self.deformation = dict(levelset_domain), definedonelements, deformation

    def __call__(self, definedonelements=None, levelset_domain=None, deformation=None, **kwargs):
        ld = dict(self.levelset_domain if levelset_domain is None else levelset_domain)
        if "order" in kwargs and kwargs["order"] is not None:
            ld["force_intorder"] = int(kwargs["order"])
        return CutDifferentialSymbol(ld, self.definedonelements if definedonelements is None else definedonelements,
                                     self.deformation if deformation is None else deformation)

    def __rmul__(self, form):
        return CutIntegrand([(form, self)])


def dCut(levelset, domain_type, order=None, subdivlvl=None, time_order=-1, levelset_domain=None, definedonelements=None, **kwargs):
    """dCut(levelset, domain_type, order = None, ...): differential symbol of a cut integral (NEG / POS / IF part)"""
    if time_order >= 0:
        raise NgsxError(-2, "space-time dCut integrators are outside the CUDA path")
    ld = dict(levelset_domain) if levelset_domain is not None else {}
    ld.update({"levelset": levelset, "domain_type": domain_type})
    if order is not None:
        ld["force_intorder"] = int(order)
    if subdivlvl:
        ld["subdivlvl"] = subdivlvl
    return CutDifferentialSymbol(ld, definedonelements, kwargs.get("deformation"))


def dx(levelset_domain=None, definedonelements=None, **kwargs):
    """dx(levelset_domain = {...}): the volume symbol restricted to a level-set domain (plain dx is NGSolve's, not part of the path)"""
    if levelset_domain is None:
        raise NgsxError(-2, "only dx(levelset_domain=...) is provided (uncut NGSolve integrators are not part of the path)")
    return CutDifferentialSymbol(levelset_domain, definedonelements, kwargs.get("deformation"))


# ------------------------------------------------------------------------------------------------------------------
# integrators
# ------------------------------------------------------------------------------------------------------------------
def _single(d: Dict[int, float], what):
    if len(d) != 1:
        raise NgsxError(-2, f"{what}: one power of h per term is supported")
    (p, c), = d.items()
    return c, p


class SymbolicCutBFI:
    """xfem/python_xfem.cpp:450-532 -> SymbolicCutBilinearFormIntegrator (xfem/symboliccutbfi.cpp:19-315)"""

    def __init__(self, lset, domain_type=NEG, force_intorder: int = -1, time_order: int = -1, subdivlvl: int = 0,
                 quad_dir_policy=OPTIMAL, form: Expr = None, VOL_or_BND=VOL, element_boundary: bool = False, skeleton: bool = False,
                 definedon=None, definedonelements: Optional[BitArray] = None, deformation=None):
        if form is None:
            raise NgsxError(-1, "form is required")
        if VOL_or_BND != VOL or element_boundary or skeleton:
            raise NgsxError(-2, "element_boundary / skeleton / BND cut integrators are outside the CUDA path")
        if time_order >= 0 or subdivlvl != 0 or definedon is not None:
            raise NgsxError(-2, "time_order / subdivlvl / definedon are outside the CUDA path")
        self.lset, self.dt, self.force_intorder, self.definedonelements = lset, int(domain_type), force_intorder, definedonelements
        self.deformation = deformation                     # bfi->SetDeformation (xfem/python_xfem.cpp:512-513); None: the mesh's
        self.xterms = None
        self.gf = getattr(form, "gf", None)                # GridFunction factor of the whole form (ngsx_elmat_h1_cf)
        from . import xfes
        if xfes.has_x_atoms(form):
            if self.gf is not None:
                raise NgsxError(-2, "a GridFunction coefficient on the compound space is outside the CUDA path")
            self._init_x(form, xfes)
            return
        terms = form.collect()
        f = _lib.FormH1()
        difforder_trial, difforder_test = 99, 99
        nit = {}
        for atoms, d in terms.items():
            c, p = _single(d, "SymbolicCutBFI")
            if atoms == ("gugv",):
                f.c_stiff, f.hpow_stiff = c, p
                difforder_trial, difforder_test = min(difforder_trial, 1), min(difforder_test, 1)
            elif atoms == ("u", "v"):
                f.c_mass, f.hpow_mass = c, p
                difforder_trial, difforder_test = 0, 0
            elif atoms in (("gun", "v"), ("gvn", "u")):
                nit[atoms] = (c, p)
            elif atoms == ("gun", "gvn"):                      # normal-normal part, e.g. of P(grad u).P(grad v), P = I - n n^T
                f.c_dndn, f.hpow_dndn = c, p
                difforder_trial, difforder_test = min(difforder_trial, 1), min(difforder_test, 1)
            else:
                raise NgsxError(-2, f"form term {atoms} is outside the CUDA path (supported: grad(u)*grad(v), u*v, "
                                    "(grad(u)*n)*v + (grad(v)*n)*u, (grad(u)*n)*(grad(v)*n) with n = LevelsetNormal()); no CPU fallback")
        if nit:
            if len(nit) != 2 or nit[("gun", "v")] != nit[("gvn", "u")]:
                raise NgsxError(-2, "the Nitsche consistency terms must appear as the symmetric pair")
            f.c_nitsche, f.hpow_nitsche = nit[("gun", "v")]
            difforder_trial, difforder_test = 0, 0           # proxies u (0) and grad u (1) both present: min = 0
        self.form = f
        self.difforders = (difforder_trial, difforder_test)

    def _init_x(self, form, xfes):
        """form on the compound space [Vh, XFESpace]: product terms for ngsx_xfem_elmat_h1 (IF elements) + the part living on the
        dofs of Vh alone, which is all that is left on uncut elements (an XFESpace has no dofs there, xfem/xFESpace.cpp:15-28)"""
        self.xterms, self.difforders = xfes.xterms_of(form)
        f = _lib.FormH1()
        self.has_std = False
        for c, p, kap, pair, bu, bv in self.xterms:
            if bu or bv:
                continue
            if self.dt == IF:
                continue                                  # no uncut elements in an interface integral
            if kap:
                raise NgsxError(-2, "a cut-ratio weight on a volume term of Vh is outside the CUDA path")
            if pair == 4 and f.c_stiff == 0.0:
                f.c_stiff, f.hpow_stiff = c, p
            elif pair == 0 and f.c_mass == 0.0:
                f.c_mass, f.hpow_mass = c, p
            else:
                raise NgsxError(-2, "volume terms on Vh: one grad(u)*grad(v) and one u*v term are supported")
            self.has_std = True
        self.form = f

    def intorder(self, fe_order: int) -> int:
        """xfem/symboliccutbfi.cpp:97-113"""
        if self.force_intorder >= 0:
            return self.force_intorder
        return 2 * fe_order - (self.difforders[0] + self.difforders[1])


class SymbolicCutLFI:
    """xfem/python_xfem.cpp:608-670 -> SymbolicCutLinearFormIntegrator (xfem/symboliccutlfi.cpp:52-158)"""

    def __init__(self, lset, domain_type=NEG, force_intorder: int = -1, time_order: int = -1, subdivlvl: int = 0,
                 quad_dir_policy=OPTIMAL, form=None, VOL_or_BND=VOL, element_boundary: bool = False, skeleton: bool = False,
                 definedon=None, definedonelements: Optional[BitArray] = None, deformation=None):
        if VOL_or_BND != VOL or element_boundary or skeleton or time_order >= 0 or subdivlvl != 0:
            raise NgsxError(-2, "this SymbolicCutLFI variant is outside the CUDA path")
        self.lset, self.dt, self.force_intorder, self.definedonelements = lset, int(domain_type), force_intorder, definedonelements
        self.deformation = deformation                     # lfi->SetDeformation (xfem/python_xfem.cpp:650-651); None: the mesh's
        # form = f * v with f a polynomial CoefficientFunction (py_tutorials/fictdom_ghostpen.py:98: coeff_f*v); on [Vh, Vhx]
        # f * (v_std + neg(v_x)) etc. (py_tutorials/nxfem.py:109-110): one scalar per test block {std, x NEG-masked, x POS-masked}
        names = {("v",): 0, ("xvn",): 1, ("xvp",): 2}
        if not isinstance(form, Expr) or not form.terms or any(p != 0 or a not in names for _, p, a in form.terms):
            raise NgsxError(-2, "SymbolicCutLFI: form must be f*v (f*(v + neg(v_x)) on [Vh, Vhx]) with a polynomial f; no CPU fallback")
        self.blocks = [0.0, 0.0, 0.0]
        for c, _, a in form.terms:
            self.blocks[names[a]] += c
        self.poly = form.poly if form.poly is not None else Poly({(0, 0, 0): 1.0})
        self.coef = self.poly * self.blocks[0]
        self.has_x = self.blocks[1] != 0.0 or self.blocks[2] != 0.0
        self.gf = getattr(form, "gf", None)                # GridFunction factor of the coefficient: f = poly * gf (ngsx_elvec_h1_cf)
        if self.gf is not None and (self.has_x or deformation is not None):
            raise NgsxError(-2, "a GridFunction coefficient is served for f*v on H1 (affine mesh); no CPU fallback")

    def intorder(self, fe_order):
        return self.force_intorder if self.force_intorder >= 0 else 2 * fe_order       # xfem/symboliccutlfi.cpp:85


class SymbolicFacetPatchBFI:
    """xfem/python_xfem.cpp:534-606 -> SymbolicFacetPatchBilinearFormIntegrator (xfem/symboliccutbfi.cpp:1202-1411)"""

    def __init__(self, form: Expr = None, force_intorder: int = -1, time_order: int = -1, skeleton: bool = True,
                 definedonelements: Optional[BitArray] = None, deformation=None):
        if skeleton:
            raise NgsxError(-2, "skeleton=True (derivative-jump ghost penalty, xfem/ghostpenalty.cpp) is outside the CUDA path; "
                                "use skeleton=False (facet-patch)")
        if time_order >= 0:
            raise NgsxError(-2, "space-time facet patches are outside the CUDA path")
        self.deformation = deformation                     # None: the mesh's (MapPatchIntegrationPoint through both curved maps)
        terms = form.collect()
        want = {("u", "v"): 1.0, ("uo", "v"): -1.0, ("u", "vo"): -1.0, ("uo", "vo"): 1.0}
        if set(terms) != set(want):
            raise NgsxError(-2, "facet-patch form must be c*h^p*(u-u.Other())*(v-v.Other())")
        c, p = _single(terms[("u", "v")], "SymbolicFacetPatchBFI")
        for a, s in want.items():
            if _single(terms[a], "SymbolicFacetPatchBFI") != (s * c, p):
                raise NgsxError(-2, "facet-patch form must be c*h^p*(u-u.Other())*(v-v.Other())")
        self.coef, self.hpow, self.definedonelements = c, p, definedonelements    # note: force_intorder is ignored by the reference too (Appendix C.2)


def SymbolicBFI(levelset_domain=None, *args, **kwargs):
    """python/__init__.py:107-192"""
    if levelset_domain is None or not isinstance(levelset_domain, dict):
        raise NgsxError(-2, "only SymbolicBFI(levelset_domain=...) is provided")
    ld = dict(levelset_domain)
    if "levelset" not in ld or "domain_type" not in ld:
        raise NgsxError(-1, "levelset_domain needs 'levelset' and 'domain_type'")
    form = kwargs.pop("form", args[0] if args else None)
    time_order = kwargs.pop("time_order", ld.get("time_order", -1))
    if time_order is not None and time_order >= 0:                  # space-time integrator (cutint/xintegration.cpp:38-58)
        from . import spacetime
        return spacetime.SymbolicSpaceTimeBFI(ld["levelset"], ld["domain_type"], form, time_order, ld.get("force_intorder", -1))
    return SymbolicCutBFI(lset=ld["levelset"], domain_type=ld["domain_type"], force_intorder=ld.get("force_intorder", -1),
                          subdivlvl=ld.get("subdivlvl", 0), quad_dir_policy=ld.get("quad_dir_policy", OPTIMAL), form=form, **kwargs)


def SymbolicLFI(levelset_domain=None, *args, **kwargs):
    """python/__init__.py:194-279"""
    if levelset_domain is None or not isinstance(levelset_domain, dict):
        raise NgsxError(-2, "only SymbolicLFI(levelset_domain=...) is provided")
    ld = dict(levelset_domain)
    form = kwargs.pop("form", args[0] if args else None)
    return SymbolicCutLFI(lset=ld["levelset"], domain_type=ld["domain_type"], force_intorder=ld.get("force_intorder", -1),
                          subdivlvl=ld.get("subdivlvl", 0), quad_dir_policy=ld.get("quad_dir_policy", OPTIMAL), form=form, **kwargs)


# ------------------------------------------------------------------------------------------------------------------
# forms: symbolic phase (graph, colouring, slot map) cached; numeric phase = one fused launch sequence per integrator
# ------------------------------------------------------------------------------------------------------------------
class _Plan:
    """colour-sorted element list + slot map of one integrator"""

    def __init__(self, els, table, color_ptr, group, slots=None):
        self.els, self.table, self.color_ptr, self.group, self.slots = els, table, color_ptr, group, slots
        self.ncolors = len(color_ptr) - 1
        self.h_color_ptr = (ctypes.c_int64 * len(color_ptr))(*color_ptr)


class _RowPlan:
    """owner-computes plan of one integrator (ngsx_rowplan): element list + per-row gather descriptors on the device"""
    ncolors = 0

    def __init__(self, els, table, handle, info):
        self.els, self.table, self.handle = els, table, handle
        self.nrows, self.npairs, self.nnz_covered, self.max_row, self.nnz_lo, self.nnz_hi = info

    def __del__(self):
        try:
            if self.handle is not None:
                _lib_().ngsx_rowplan_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


def _packed_rows(plan, nd):
    """row-slot map of a plan (device pointer) and the packed rows buffer [nslots][nd rounded up to even], allocated once"""
    if getattr(plan, "rows", None) is None:
        p, ns = ctypes.c_void_p(), ctypes.c_int64(0)
        check(_lib_().ngsx_rowplan_rowslots(plan.handle, ctypes.byref(p), ctypes.byref(ns), _stream()))
        plan.slots = p
        plan.rows = torch.zeros(max(1, ns.value) * (nd + (nd & 1)), dtype=torch.float64, device=plan.els.device)
    return plan.slots, plan.rows


def color_list(table: torch.Tensor, ndof: int, group: int):
    """ngsx_color_groups on the host copy of a dof table [n][nd]; returns (perm (device), color_ptr list)"""
    n, nd = table.shape
    h_table = table.cpu().contiguous()
    perm = torch.empty(n, dtype=torch.int64)
    cap = 1024
    cptr = torch.zeros(cap + 1, dtype=torch.int64)
    nc = ctypes.c_int32(0)
    check(_lib_().ngsx_color_groups(n, nd, h_table.data_ptr(), ndof, group, perm.data_ptr(), cptr.data_ptr(), cap, ctypes.byref(nc)))
    return perm.to(table.device), cptr[: nc.value + 1].tolist()



class _XPlan:
    """compound integrator: its product terms, the row plan of its Vh part on the uncut elements, and the gather plan + block
    buffer over [dofs of Vh | x-dofs] of the IF elements that all compound integrators of the form share"""

    def __init__(self, shared, xterms, plan_a):
        self.shared, self.xterms, self.plan_a = shared, xterms, plan_a
        self.els, self.handle, self.mats = shared.els, shared.handle, shared.mats

class BilinearForm:
    """T_BilinearForm<double> stand-in: `a += integrator`, `a.Assemble()`, `a.mat` (CSR on the device).
    The sparsity is the full element (and, with dgjumps, facet-patch) graph like NGSolve's GetGraph."""
    group_size = 6
    assembly = "rows"          # "rows": owner-computes row gather (default); "colored": colour-ordered scatter-add

    def __new__(cls, space=None, *args, **kwargs):
        from .spacetime import SpaceTimeBilinearForm, SpaceTimeFESpace
        if cls is BilinearForm and isinstance(space, SpaceTimeFESpace):      # tensor-product space: blocks from ngsx_st_elmat_h1
            return SpaceTimeBilinearForm(space)
        return super().__new__(cls)

    def __init__(self, space: H1, symmetric: bool = False, check_unused: bool = False, **_):
        self.space = space
        # symmetric=True: SparseMatrixSymmetric storage -- the CSR keeps the lower triangle (columns <= row) only and the gather skips
        # the entries above it (AddElementMatrixSymmetric, NGS/linalg/sparsematrix_impl.hpp:605-690); values are bitwise the lower
        # triangle of the unsymmetric assembly.  Served by the fused H1 row path (orders 1-2, affine mesh, single GPU).
        self.symmetric = bool(symmetric)
        self.integrators: List = []
        self.mat: Optional[CSRMatrix] = None
        self._plans: Dict[int, _Plan] = {}
        self._xshared = None
        self._graph_key = None
        self._err = torch.zeros((1, 2), dtype=torch.int64, device=space.mesh.device)     # one {count, first} per integrator
        self._err_armed = False
        self.row_range = None      # (lo, hi): this rank owns CSR rows [lo, hi) (multi-GPU owner-computes partition)
        self.row_owner = None      # (owner int32 [ndof] on the device, me): this rank owns the rows with owner == me
        self._h1 = getattr(space, "std", space)        # the H1 component (compound space [Vh, XFESpace]: its first component)

    def __iadd__(self, integ):
        if isinstance(integ, CutIntegrand):                            # a += form * dCut(...) [+ form2 * dCut(...)]
            for form, sym in integ.parts:
                self += SymbolicBFI(levelset_domain=sym.levelset_domain, form=form, definedonelements=sym.definedonelements,
                                    deformation=sym.deformation)
            return self
        if not isinstance(integ, (SymbolicCutBFI, SymbolicFacetPatchBFI)):
            raise NgsxError(-2, "only SymbolicCutBFI / SymbolicFacetPatchBFI integrators are on the CUDA path")
        self.integrators.append(integ)
        return self

    # --- which elements / facets couple (overridden by RestrictedBilinearForm) ---
    def _graph_elements(self) -> Optional[torch.Tensor]:
        return None            # all

    def _graph_facets(self) -> Optional[torch.Tensor]:
        if not self.space.dgjumps:
            return torch.zeros(0, dtype=torch.int64, device=self.space.mesh.device)
        _, f2e = self.space.mesh.facets()
        return torch.nonzero(f2e[:, 1] >= 0).reshape(-1)

    def _build_graph(self):
        """RestrictedBilinearForm::GetGraph (utils/restrictedblf.cpp:22-138) on the device"""
        L = _lib_()
        fes = self.space
        g = ctypes.c_void_p()
        check(L.ngsx_csr_graph_create(fes.ndof, ctypes.byref(g)))
        try:
            if self.symmetric:
                check(L.ngsx_csr_graph_set_lower(g, 1))
            els = self._graph_elements()
            h1 = self._h1
            tab = h1.el2dof if els is None else h1.el2dof[els].contiguous()
            check(L.ngsx_csr_graph_add_table(g, tab.shape[0], h1.nd, _ptr(tab), _stream()))
            if h1 is not fes:                                      # compound space: couplings of the x-dofs on the IF elements
                ctab = fes.compound_table()
                if ctab.shape[0]:
                    check(L.ngsx_csr_graph_add_table(g, ctab.shape[0], ctab.shape[1], _ptr(ctab), _stream()))
            fac = self._graph_facets()
            if fac is not None and fac.numel() > 0:
                _, f2e = fes.mesh.facets()
                pe = f2e[fac].long()
                ptab = torch.cat([h1.el2dof[pe[:, 0]], h1.el2dof[pe[:, 1]]], dim=1).contiguous()
                check(L.ngsx_csr_graph_add_table(g, ptab.shape[0], 2 * h1.nd, _ptr(ptab), _stream()))
            nnz = ctypes.c_int64(0)
            check(L.ngsx_csr_graph_finalize(g, ctypes.byref(nnz), _stream()))
            if nnz.value >= 2 ** 31:
                raise NgsxError(-5, "more than 2^31 non-zeros: int32 slot map exhausted")
            dev = fes.mesh.device
            rowptr = torch.empty(fes.ndof + 1, dtype=torch.int64, device=dev)
            colidx = torch.empty(nnz.value, dtype=torch.int32, device=dev)
            check(L.ngsx_csr_graph_export(g, _ptr(rowptr), _ptr(colidx), _stream()))
            torch.cuda.current_stream().synchronize()
        finally:
            L.ngsx_csr_graph_destroy(g)
        self.mat = CSRMatrix(fes.ndof, rowptr, colidx, torch.zeros(nnz.value, dtype=torch.float64, device=dev))

    def _active_elements(self, integ: SymbolicCutBFI) -> torch.Tensor:
        """elements on which the reference's CalcElementMatrixAdd gets a non-null rule (+ definedonelements)"""
        ci = CutInfo(self.space.mesh, integ.lset)
        cdt = {NEG: HASNEG, POS: HASPOS, IF: CDOM_IF}[integ.dt]
        ba = ci.GetElementsOfType(cdt)
        if integ.definedonelements is not None:
            ba = ba & integ.definedonelements
        return ba.nonzero()

    def _listed_elements(self, integ: SymbolicCutBFI) -> torch.Tensor:
        """elements the integrator may ever touch: the graph's elements restricted by definedonelements.  Independent of the
        current level-set values, so the plan survives a level-set update (the kernel re-classifies every listed element)."""
        els = self._graph_elements()
        if integ.definedonelements is not None:
            de = integ.definedonelements.to_bool()
            els = torch.nonzero(de).reshape(-1) if els is None else els[de[els]]
        if els is None:
            els = torch.arange(self.space.mesh.ne, device=self.space.mesh.device)
        if integ.xterms is not None:
            # the IF elements go through the compound kernel, and this plan is tied to the XFESpace's cut state anyway (it is
            # rebuilt with it): list only the uncut elements of the integrator's own domain instead of re-classifying all
            cdt = {NEG: CDOM_NEG, POS: CDOM_POS}.get(integ.dt)
            if cdt is None:
                return els[:0]
            els = els[self.space.x.cutinfo.GetElementsOfType(cdt).to_bool()[els]]
        return els

    def _x_plan(self, integ: SymbolicCutBFI):
        """compound integrator: row plan of the Vh part on the uncut elements + row plan over [dofs of Vh | x-dofs] of the IF elements"""
        fes, L = self.space, _lib_()
        if self._h1 is fes:
            raise NgsxError(-1, "neg / pos operators need a BilinearForm on FESpace([Vh, XFESpace(Vh)])")
        plan_a = self._row_plan(integ) if (integ.dt != IF and integ.has_std) else None
        if self._xshared is None:                                  # one gather plan + one block buffer for all compound integrators
            ctab = fes.compound_table()
            h = ctypes.c_void_p()
            lo, hi = self.row_range if self.row_range is not None else (0, fes.ndof)
            check(L.ngsx_rowplan_create(ctab.shape[0], ctab.shape[1], _ptr(ctab), fes.ndof, _ptr(self.mat.rowptr), _ptr(self.mat.colidx),
                                        lo, hi, ctypes.byref(h), _stream()))
            info = (ctypes.c_int64 * 6)()
            check(L.ngsx_rowplan_info(h, info))
            shared = _RowPlan(fes.x.if_els, ctab, h, tuple(info))
            shared.slots, shared.mats = _packed_rows(shared, ctab.shape[1])     # packed rows of the compound blocks (2 nd is even)
            self._xshared = shared
        from . import xfes
        plan_b = _XPlan(self._xshared, xfes.make_xterms(integ.xterms), plan_a)
        return plan_b

    def _row_plan(self, integ: SymbolicCutBFI) -> _RowPlan:
        fes, L = self._h1, _lib_()
        els = self._listed_elements(integ).to(torch.int32).contiguous()
        table = fes.el2dof[els.long()].contiguous()
        h = ctypes.c_void_p()
        lo, hi = self.row_range if self.row_range is not None else (0, self.space.ndof)
        if self.symmetric:
            if self.row_owner is not None:
                raise NgsxError(-2, "symmetric storage with an owner partition is outside the CUDA path")
            check(L.ngsx_rowplan_create_lower(els.numel(), fes.nd, _ptr(table), self.space.ndof, _ptr(self.mat.rowptr), _ptr(self.mat.colidx),
                                              lo, hi, ctypes.byref(h), _stream()))
        elif self.row_owner is not None:
            check(L.ngsx_rowplan_create_owned(els.numel(), fes.nd, _ptr(table), self.space.ndof, _ptr(self.mat.rowptr), _ptr(self.mat.colidx),
                                              _ptr(self.row_owner[0]), int(self.row_owner[1]), ctypes.byref(h), _stream()))
        else:
            check(L.ngsx_rowplan_create(els.numel(), fes.nd, _ptr(table), self.space.ndof, _ptr(self.mat.rowptr), _ptr(self.mat.colidx),
                                        lo, hi, ctypes.byref(h), _stream()))
        info = (ctypes.c_int64 * 6)()
        check(L.ngsx_rowplan_info(h, info))
        return _RowPlan(els, table, h, tuple(info))

    def _plan_for(self, k: int, integ) -> _Plan:
        if k in self._plans:
            return self._plans[k]
        fes, L = self.space, _lib_()
        if isinstance(integ, SymbolicCutBFI) and integ.xterms is not None:
            plan = self._x_plan(integ)
        elif isinstance(integ, SymbolicCutBFI) and self.assembly == "rows":
            try:
                plan = self._row_plan(integ)
            except NgsxError as exc:
                if exc.status != -2 or self.symmetric:
                    raise
                import warnings
                warnings.warn("ngsxfem_b200: the owner-computes row plan does not cover this pattern (" + str(exc) +
                              "); falling back to the coloured scatter, which is several times slower", RuntimeWarning)
                self.assembly = "colored"              # a CSR row longer than 256 entries: coloured scatter instead
                return self._plan_for(k, integ)
        elif isinstance(integ, SymbolicCutBFI):
            els = self._listed_elements(integ).long()                  # independent of the level-set values: the kernel classifies
            table = fes.el2dof[els].contiguous()
            perm, cptr = color_list(table, fes.ndof, self.group_size)
            els_s = els[perm].to(torch.int32).contiguous()
            table_s = table[perm].contiguous()
            slots = torch.empty((els_s.numel(), fes.nd * fes.nd), dtype=torch.int32, device=fes.mesh.device)
            check(L.ngsx_slot_map(els_s.numel(), fes.nd, _ptr(table_s), _ptr(self.mat.rowptr), _ptr(self.mat.colidx), _ptr(slots), _stream()))
            if els_s.numel() and int(slots.min().item()) < 0:
                raise NgsxError(-1, "an element coupling is missing from the sparsity pattern (element_restriction too small?)")
            plan = _Plan(els_s, table_s, cptr, self.group_size, slots)
        else:
            plan = self._facetpatch_plan(integ)
        self._plans[k] = plan
        return plan

    def _facetpatch_plan(self, integ: SymbolicFacetPatchBFI):
        """symbolic phase of the ghost-penalty operator (NGS/comp/bilinearform.cpp:1800-1853): patches = inner facets (restricted
        by definedonelements, which is a FACET BitArray for this integrator, xfem/python_xfem.cpp:534-606), merged dof list per
        patch (ngsx_facetpatch_table_h1) and an owner-computes row plan over it"""
        fes, L = self.space, _lib_()
        mesh = fes.mesh
        _, f2e = mesh.facets()
        inner = f2e[:, 1] >= 0
        if integ.definedonelements is not None:
            inner = inner & integ.definedonelements.to_bool()
        fac = torch.nonzero(inner).reshape(-1)
        pe = f2e[fac].contiguous()                                     # [np][2] int32, element 0 = lower number
        npatch = pe.shape[0]
        if (mesh.dim, fes.order) not in ((2, 1), (2, 2), (3, 1), (3, 2), (2, 3)):
            raise NgsxError(-2, "SymbolicFacetPatchBFI: H1 order 1..2 (and order 3 on triangles) is on the CUDA path -- a P3 tet patch has 30 "
                                "merged dofs, the gather's descriptors end at 26; no CPU fallback")
        nfd = {(2, 1): 2, (2, 2): 3, (3, 1): 3, (3, 2): 6, (2, 3): 4}[(mesh.dim, fes.order)]          # dofs on the shared facet
        ndp = 2 * fes.nd - nfd
        table = torch.empty((npatch, ndp), dtype=torch.int32, device=mesh.device)
        cidx = torch.empty((npatch, 2 * fes.nd), dtype=torch.uint8, device=mesh.device)
        check(L.ngsx_facetpatch_table_h1(npatch, _ptr(pe), fes.nd, _ptr(fes.el2dof), ndp, _ptr(table), _ptr(cidx), _stream()))
        h = ctypes.c_void_p()
        lo, hi = self.row_range if self.row_range is not None else (0, fes.ndof)
        check(L.ngsx_rowplan_create(npatch, ndp, _ptr(table), fes.ndof, _ptr(self.mat.rowptr), _ptr(self.mat.colidx),
                                    lo, hi, ctypes.byref(h), _stream()))
        info = (ctypes.c_int64 * 6)()
        check(L.ngsx_rowplan_info(h, info))
        plan = _RowPlan(pe, table, h, tuple(info))
        plan.cidx, plan.ndp = cidx, ndp
        return plan

    def Assemble(self, reallocate: bool = False, gather_after=None):
        """S_BilinearForm::DoAssemble (NGS/comp/bilinearform.cpp:929-1575): SetZero, then every integrator colour by colour.
        gather_after: a recorded torch.cuda.Event of a side stream -- the row gather of the (first) H1 row-plan integrator starts
        only after it (ngsx_set_gather_wait_event), the geometry / quadrature stage before it overlaps the side work"""
        L = _lib_()
        fes = self.space
        xver = getattr(getattr(fes, "x", None), "version", None)      # XFESpace.Update() changes the x-dofs: graph + plans are stale
        if self.mat is None or reallocate or xver != self._graph_key:
            self._build_graph()
            self._plans.clear()
            self._xshared = None
            self._graph_key = xver
        if fes.mesh.deformation is not None or any(getattr(i, "deformation", None) is not None for i in self.integrators):
            for integ in self.integrators:                         # what the deformed path covers
                if isinstance(integ, SymbolicFacetPatchBFI):
                    continue
                if integ.xterms is None and self.assembly != "rows":
                    raise NgsxError(-2, "on a deformed mesh the row gather is the CUDA path (the coloured scatter is undeformed only); "
                                        "no CPU fallback")
        plans = [self._plan_for(k, integ) for k, integ in enumerate(self.integrators)]
        # a single row plan that covers every stored entry overwrites its rows: no SetZero pass, each value written once
        # (with a row range: the rank's rows are one contiguous CSR span, the other rows belong to other ranks and stay untouched)
        overwrite = len(plans) == 1 and isinstance(self.integrators[0], SymbolicCutBFI) and self.integrators[0].xterms is None and \
            isinstance(plans[0], _RowPlan) and \
            (self.row_owner is not None or
             plans[0].nnz_covered == (self.mat.nnz if self.row_range is None else plans[0].nnz_hi - plans[0].nnz_lo))
        if not overwrite:
            self.mat.vals.zero_()
        if self._err.shape[0] < len(plans):
            self._err = torch.zeros((len(plans), 2), dtype=torch.int64, device=fes.mesh.device)
        mesh_c = ctypes.byref(fes.mesh.c_struct())
        x_blocks = 0                                               # compound integrators whose IF blocks sit in the shared buffer
        for k, integ in enumerate(self.integrators):
            plan = plans[k]
            if self.symmetric and not (isinstance(plan, _RowPlan) and isinstance(integ, SymbolicCutBFI) and integ.xterms is None and
                                       getattr(integ, "gf", None) is None and
                                       self._deformation_of(integ) is None and fes.order < 3 and self._h1 is fes):
                raise NgsxError(-2, "symmetric=True is served by the fused H1 row path (SymbolicCutBFI, orders 1-2, affine mesh); no CPU fallback")
            plain_rows = isinstance(plan, _RowPlan) and not isinstance(integ, SymbolicFacetPatchBFI) and integ.xterms is None and \
                self._deformation_of(integ) is None and fes.order < 3 and getattr(integ, "gf", None) is None
            if gather_after is not None and not plain_rows:
                # only the fused H1 row path can hold back its gather alone (ngsx_set_gather_wait_event); every other producer
                # honours the event by waiting for it on the stream before it starts
                torch.cuda.current_stream().wait_event(gather_after)
                gather_after = None
            if isinstance(integ, SymbolicFacetPatchBFI):
                self._assemble_facetpatch(integ, plan)
            elif integ.xterms is not None:
                x_blocks += self._assemble_x(integ, plan, k, x_blocks)
            elif isinstance(plan, _RowPlan) and getattr(integ, "gf", None) is not None:
                # a GridFunction factor in the form: the coefficient varies inside the element, every listed element goes through the
                # quadrature kernel (dense blocks), then the owner-computes gather
                cfs = integ.gf.space
                if self._deformation_of(integ) is not None or fes.order >= 3 or cfs.mesh is not fes.mesh:
                    raise NgsxError(-2, "a GridFunction coefficient is served on affine meshes for H1 order 1-2 (same mesh); no CPU fallback")
                lv = _lset_vec(integ.lset)
                n = plan.els.numel()
                if getattr(plan, "mats", None) is None:
                    plan.mats = torch.empty((n, fes.nd * fes.nd), dtype=torch.float64, device=fes.mesh.device)
                check(L.ngsx_elmat_h1_cf(mesh_c, _ptr(lv), fes.order, integ.dt, integ.intorder(fes.order), ctypes.byref(integ.form), cfs.order,
                                         _ptr(cfs.el2dof), _ptr(integ.gf.vec), n, _ptr(plan.els), _ptr(plan.mats), None, _stream()))
                check(L.ngsx_assemble_csr_rows_mats(plan.handle, _ptr(plan.mats), 0 if overwrite else 1, _ptr(self.mat.vals), _stream()))
            elif isinstance(plan, _RowPlan) and (self._deformation_of(integ) is not None or fes.order >= 3):
                # deformed elements (row f-4: per-point geometry) and H1 order 3 (orientation-dependent edge functions: no shared
                # constant-rule table): dense blocks of the listed elements, then the owner-computes gather (each CSR value
                # written once, plan order, no atomics)
                lv = _lset_vec(integ.lset)
                n = plan.els.numel()
                if self._deformation_of(integ) is not None:
                    if fes.order >= 3:
                        raise NgsxError(-2, "H1 order 3 on a deformed mesh is outside the CUDA path; no CPU fallback")
                    # packed block rows: the kernel writes row i of block k where the gather's schedule reads it (ngsx_rowplan_rowslots)
                    slots, rows = _packed_rows(plan, fes.nd)
                    defo = fes.mesh.c_deformation(self._deformation_of(integ))
                    L.ngsx_set_packed_output(slots)
                    check(L.ngsx_elmat_h1_deformed(mesh_c, ctypes.byref(defo), _ptr(lv), fes.order, integ.dt, integ.intorder(fes.order),
                                                   ctypes.byref(integ.form), n, _ptr(plan.els), _ptr(rows), None, _ptr(self._err[k]), _stream()))
                    self._err_armed = True
                    check(L.ngsx_assemble_csr_rows_packed(plan.handle, _ptr(rows), 0 if overwrite else 1, _ptr(self.mat.vals), _stream()))
                else:
                    if getattr(plan, "mats", None) is None:
                        plan.mats = torch.empty((n, fes.nd * fes.nd), dtype=torch.float64, device=fes.mesh.device)
                    # order 3: k_elmat3_fast (sign-flipped constant-rule tables) + k_elmat3 (DMMA contraction on the cut elements)
                    check(L.ngsx_elmat_h1(mesh_c, _ptr(lv), fes.order, integ.dt, integ.intorder(fes.order), ctypes.byref(integ.form), n,
                                          _ptr(plan.els), _ptr(plan.mats), None, _stream()))
                    check(L.ngsx_assemble_csr_rows_mats(plan.handle, _ptr(plan.mats), 0 if overwrite else 1, _ptr(self.mat.vals), _stream()))
            elif isinstance(plan, _RowPlan):
                lv = _lset_vec(integ.lset)
                if gather_after is not None:
                    L.ngsx_set_gather_wait_event(ctypes.c_void_p(gather_after.cuda_event))
                    gather_after = None
                check(L.ngsx_assemble_csr_rows_h1(mesh_c, _ptr(lv), fes.order, integ.dt, integ.intorder(fes.order), ctypes.byref(integ.form),
                                                  plan.els.numel(), _ptr(plan.els), plan.handle, 0 if overwrite else 1,
                                                  _ptr(self.mat.vals), _ptr(self._err[k]), _stream()))
                self._err_armed = True
            elif isinstance(integ, SymbolicCutBFI):
                if getattr(integ, "gf", None) is not None:
                    raise NgsxError(-2, "a GridFunction coefficient is served by the row gather (assembly = 'rows'); no CPU fallback")
                lv = _lset_vec(integ.lset)
                check(L.ngsx_assemble_csr_h1(mesh_c, _ptr(lv), fes.order, integ.dt, integ.intorder(fes.order), ctypes.byref(integ.form),
                                             plan.els.numel(), _ptr(plan.els), _ptr(plan.slots), plan.ncolors, plan.h_color_ptr,
                                             plan.group, _ptr(self.mat.vals), _stream()))
        if x_blocks:                                               # one gather for the summed blocks of all compound integrators
            check(L.ngsx_assemble_csr_rows_packed(self._xshared.handle, _ptr(self._xshared.mats), 1, _ptr(self.mat.vals), _stream()))
        return self

    def _deformation_of(self, integ):
        """the integrator's own deformation, else the mesh's (NGS/comp/bilinearform.cpp:1235)"""
        d = getattr(integ, "deformation", None)
        return d if d is not None else self.space.mesh.deformation

    def check_degenerate(self, err=None):
        """raise if the last Assemble met a level set the reference throws on; synchronises.
        err: host copy of degenerate_flags() the caller downloaded itself (pipelined callers: no synchronisation here)"""
        if err is None:
            if not self._err_armed:
                return
            self._err_armed = False
            err = self._err.tolist()
            self._check_cut_capacity()
        else:
            err = err.tolist()
        n, first = sum(r[0] for r in err), min(r[1] for r in err)
        if n:
            raise NgsxError(-4, f"{n} element(s) carry a level set the reference rejects (first: element {first}): "
                                "'You tried to cut a simplex with a plain geometry lset function'")

    def _check_cut_capacity(self):
        """the cut-element buffer of a row plan is sized on its first numeric call; a level set that needs more is reported here
        (the excess elements of that Assemble were skipped): the buffer is grown and the caller assembles again"""
        L = _lib_()
        for plan in self._plans.values():
            for pl in (plan, getattr(plan, "plan_a", None)):
                if isinstance(pl, _RowPlan) and pl.handle is not None:
                    cnt, cap = ctypes.c_int64(0), ctypes.c_int64(0)
                    check(L.ngsx_rowplan_cut_usage(pl.handle, ctypes.byref(cnt), ctypes.byref(cap), _stream()))
                    if cap.value and cnt.value > cap.value:
                        check(L.ngsx_rowplan_reserve_cut(pl.handle, 2 * cnt.value))
                        raise NgsxError(-5, f"{cnt.value} cut elements exceeded the plan's buffer ({cap.value}); it has been grown: Assemble again")

    def degenerate_flags(self):
        """device tensor [{count, first element}] per integrator of the last Assemble, for callers that download it on their
        own stream together with the values and hand the host copy to check_degenerate(err)"""
        self._err_armed = False
        return self._err

    def _assemble_x(self, integ, plan, k, x_blocks):
        """compound integrator: Vh part on the uncut elements through the H1 row gather; the (2 nd)^2 blocks of the IF elements
        (ngsx_xfem_elmat_h1) are summed over the form's compound integrators in one buffer that Assemble gathers once"""
        L = _lib_()
        fes, h1 = self.space, self._h1
        lv = _lset_vec(integ.lset)
        io = integ.intorder(h1.order)
        mesh_c = ctypes.byref(fes.mesh.c_struct())
        gfd = self._deformation_of(integ)
        defo = fes.mesh.c_deformation(gfd) if gfd is not None else None
        if plan.plan_a is not None:
            pa = plan.plan_a
            if defo is not None:                                       # deformed mesh: dense blocks with per-point geometry, packed rows
                slots, rows = _packed_rows(pa, h1.nd)
                L.ngsx_set_packed_output(slots)
                check(L.ngsx_elmat_h1_deformed(mesh_c, ctypes.byref(defo), _ptr(lv), h1.order, integ.dt, io, ctypes.byref(integ.form),
                                               pa.els.numel(), _ptr(pa.els), _ptr(rows), None, _ptr(self._err[k]), _stream()))
                check(L.ngsx_assemble_csr_rows_packed(pa.handle, _ptr(rows), 1, _ptr(self.mat.vals), _stream()))
            else:
                check(L.ngsx_assemble_csr_rows_h1(mesh_c, _ptr(lv), h1.order, integ.dt, io, ctypes.byref(integ.form), pa.els.numel(), _ptr(pa.els),
                                                  pa.handle, 1, _ptr(self.mat.vals), _ptr(self._err[k]), _stream()))
            self._err_armed = True
        n = plan.els.numel()
        if n:
            L.ngsx_set_packed_output(plan.shared.slots)
            if defo is not None:
                check(L.ngsx_xfem_elmat_h1_deformed(mesh_c, ctypes.byref(defo), _ptr(lv), h1.order, integ.dt, io, len(integ.xterms), plan.xterms,
                                                    n, _ptr(plan.els), _ptr(fes.x.xsign), _ptr(fes.x.cutinfo.ratio), 1 if x_blocks else 0,
                                                    _ptr(plan.mats), _stream()))
            else:
                check(L.ngsx_xfem_elmat_h1(mesh_c, _ptr(lv), h1.order, integ.dt, io, len(integ.xterms), plan.xterms, n, _ptr(plan.els),
                                           _ptr(fes.x.xsign), _ptr(fes.x.cutinfo.ratio), 1 if x_blocks else 0, _ptr(plan.mats), _stream()))
            return 1
        return 0

    def _assemble_facetpatch(self, integ, plan):
        """numeric phase: merged patch blocks (ngsx_facetpatch_cmat_h1), then the owner-computes gather adds them to the rows
        they touch (ngsx_assemble_csr_rows_mats) -- plan order, no atomics, bitwise reproducible"""
        L = _lib_()
        fes = self.space
        npatch = plan.els.shape[0]
        if npatch == 0:
            return
        slots, rows = _packed_rows(plan, plan.ndp)                     # the kernel writes block rows where the gather's schedule reads them
        L.ngsx_set_packed_output(slots)
        gfd = self._deformation_of(integ)
        if gfd is not None:                                            # deformed mesh: Newton pull-back through both curved maps
            defo = fes.mesh.c_deformation(gfd)
            check(L.ngsx_facetpatch_cmat_h1_deformed(ctypes.byref(fes.mesh.c_struct()), ctypes.byref(defo), fes.order, integ.coef, integ.hpow,
                                                     npatch, _ptr(plan.els), _ptr(plan.cidx), plan.ndp, _ptr(rows), _stream()))
        else:
            check(L.ngsx_facetpatch_cmat_h1(ctypes.byref(fes.mesh.c_struct()), fes.order, integ.coef, integ.hpow, npatch,
                                            _ptr(plan.els), _ptr(plan.cidx), plan.ndp, _ptr(rows), _stream()))
        check(L.ngsx_assemble_csr_rows_packed(plan.handle, _ptr(rows), 1, _ptr(self.mat.vals), _stream()))

    def elmats(self, k: int = 0):
        """dense element matrices of integrator k for its active elements (the per-element view CalcElementMatrixAdd would serve)"""
        integ, fes, L = self.integrators[k], self.space, _lib_()
        if getattr(integ, "gf", None) is not None:
            raise NgsxError(-2, "elmats(): integrators with a GridFunction coefficient are served through Assemble()")
        els = self._active_elements(integ).to(torch.int32).contiguous()
        out = torch.empty((els.numel(), fes.nd, fes.nd), dtype=torch.float64, device=fes.mesh.device)
        npts = torch.empty(els.numel(), dtype=torch.int32, device=fes.mesh.device)
        if self._deformation_of(integ) is not None:
            defo = fes.mesh.c_deformation(self._deformation_of(integ))
            check(L.ngsx_elmat_h1_deformed(ctypes.byref(fes.mesh.c_struct()), ctypes.byref(defo), _ptr(_lset_vec(integ.lset)), fes.order, integ.dt,
                                           integ.intorder(fes.order), ctypes.byref(integ.form), els.numel(), _ptr(els), _ptr(out), _ptr(npts),
                                           None, _stream()))
            return els, out, npts
        check(L.ngsx_elmat_h1(ctypes.byref(fes.mesh.c_struct()), _ptr(_lset_vec(integ.lset)), fes.order, integ.dt, integ.intorder(fes.order),
                              ctypes.byref(integ.form), els.numel(), _ptr(els), _ptr(out), _ptr(npts), _stream()))
        return els, out, npts


class RestrictedBilinearForm(BilinearForm):
    """utils/python_utils.cpp:86-153 / utils/restrictedblf.cpp: sparsity restricted to marked elements / facets"""

    def __init__(self, space: H1, name: str = "bfa", element_restriction: Optional[BitArray] = None,
                 facet_restriction: Optional[BitArray] = None, check_unused: bool = True, **flags):
        super().__init__(space, **flags)
        self.element_restriction, self.facet_restriction = element_restriction, facet_restriction

    def _graph_elements(self):
        return None if self.element_restriction is None else self.element_restriction.nonzero()

    def _graph_facets(self):
        if self.facet_restriction is None:
            return super()._graph_facets()
        _, f2e = self.space.mesh.facets()
        return torch.nonzero(self.facet_restriction.to_bool() & (f2e[:, 1] >= 0)).reshape(-1)


class LinearForm:
    group_size = 6

    def __init__(self, space):
        from .xfes import CompoundFESpace
        self.space = space
        self.compound = isinstance(space, CompoundFESpace)
        self._h1 = space.std if self.compound else space
        self.integrators: List[SymbolicCutLFI] = []
        self.vec = torch.zeros(space.ndof, dtype=torch.float64, device=space.mesh.device)
        self._xplans: Dict = {}          # ("h1", k): per-dof gather plan of integrator k; k: x-dof plans (tied to the XFESpace version)
        self._xver = None

    def __iadd__(self, integ):
        if isinstance(integ, CutIntegrand):                            # f += form * dCut(...)
            for form, sym in integ.parts:
                self += SymbolicLFI(levelset_domain=sym.levelset_domain, form=form, definedonelements=sym.definedonelements,
                                    deformation=sym.deformation)
            return self
        if not isinstance(integ, SymbolicCutLFI):
            raise NgsxError(-2, "only SymbolicCutLFI integrators are on the CUDA path")
        if integ.has_x and not self.compound:
            raise NgsxError(-1, "neg / pos test functions need a LinearForm on FESpace([Vh, XFESpace])")
        self.integrators.append(integ)
        return self

    def _listed_elements(self, integ) -> torch.Tensor:
        """elements the integrator may ever touch (all, or definedonelements): independent of the level-set values, so the plan
        survives a level-set update -- the kernels classify every listed element on every Assemble, like the reference's loop"""
        mesh = self._h1.mesh
        if integ.definedonelements is not None:
            return torch.nonzero(integ.definedonelements.to_bool()).reshape(-1).to(torch.int32).contiguous()
        return torch.arange(mesh.ne, dtype=torch.int32, device=mesh.device)

    def _gather_plan(self, key, integ):
        """per-dof gather plan of AddElementVector (NGS/comp/linearform.cpp:209-243 without atomics): for every dof the entries of
        the flattened element vectors that belong to it, in (element, local dof) order"""
        if key not in self._xplans:
            fes = self._h1
            els = self._listed_elements(integ)
            n, nd = els.numel(), fes.nd
            dof = fes.el2dof[els.long()].reshape(-1).long()
            src = torch.arange(n * nd, device=dof.device)
            order = torch.argsort(dof * max(n * nd, 1) + src)
            dofptr = torch.zeros(fes.ndof + 1, dtype=torch.int64, device=dof.device)
            dofptr[1:] = torch.cumsum(torch.bincount(dof, minlength=fes.ndof), 0)
            self._xplans[key] = (els, dofptr, src[order].contiguous(), torch.empty((n, nd), dtype=torch.float64, device=dof.device))
        return self._xplans[key]

    def Assemble(self):
        """NGS/comp/linearform.cpp:209-243: element vectors of all listed elements (ngsx_elvec_h1 classifies them), then the
        per-dof gather (ngsx_gather_vec: fixed order, no atomics, bitwise reproducible)"""
        L = _lib_()
        fes = self._h1
        xver = getattr(getattr(self.space, "x", None), "version", None)
        if self.vec.numel() != self.space.ndof or xver != self._xver:       # XFESpace.Update changed the x-dofs
            if self.vec.numel() != self.space.ndof:
                self.vec = torch.zeros(self.space.ndof, dtype=torch.float64, device=fes.mesh.device)
            self._xplans = {k: v for k, v in self._xplans.items() if isinstance(k, tuple)}      # keep the H1 gather plans
            self._xver = xver
        self.vec.zero_()
        for k, integ in enumerate(self.integrators):
            gfd = integ.deformation if getattr(integ, "deformation", None) is not None else fes.mesh.deformation
            if getattr(integ, "gf", None) is not None and (gfd is not None or fes.order > 2 or integ.gf.space.mesh is not fes.mesh):
                raise NgsxError(-2, "a GridFunction coefficient is served on affine meshes for H1 order 1-2 (same mesh); no CPU fallback")
            if gfd is not None:
                self._assemble_deformed(integ, k, gfd)
                if integ.has_x:
                    self._assemble_x(integ, k, gfd)
                continue
            if integ.blocks[0] != 0.0:
                els, dofptr, pairs, elvecs = self._gather_plan(("h1", k), integ)
                if els.numel():
                    poly = _poly_of(integ.coef)
                    if getattr(integ, "gf", None) is not None:
                        cfs = integ.gf.space
                        check(L.ngsx_elvec_h1_cf(ctypes.byref(fes.mesh.c_struct()), _ptr(_lset_vec(integ.lset)), fes.order, integ.dt,
                                                 integ.intorder(fes.order), ctypes.byref(poly), cfs.order, _ptr(cfs.el2dof), _ptr(integ.gf.vec),
                                                 els.numel(), _ptr(els), _ptr(elvecs), None, _stream()))
                    else:
                        check(L.ngsx_elvec_h1(ctypes.byref(fes.mesh.c_struct()), _ptr(_lset_vec(integ.lset)), fes.order, integ.dt,
                                              integ.intorder(fes.order), ctypes.byref(poly), els.numel(), _ptr(els), _ptr(elvecs), None, _stream()))
                    check(L.ngsx_gather_vec(fes.ndof, _ptr(dofptr), _ptr(pairs), _ptr(elvecs), 1.0, 1, _ptr(self.vec), _stream()))
            if integ.has_x:
                self._assemble_x(integ, k)
        return self

    def _assemble_deformed(self, integ, k, gfd):
        """deformed mesh (row f-4): dense element vectors with per-point geometry (ngsx_elvec_h1_deformed), then the per-dof gather"""
        L = _lib_()
        fes = self._h1
        if integ.blocks[0] == 0.0:
            return
        els, dofptr, pairs, elvecs = self._gather_plan(("h1", k), integ)
        if els.numel() == 0:
            return
        poly = _poly_of(integ.coef)
        defo = fes.mesh.c_deformation(gfd)
        check(L.ngsx_elvec_h1_deformed(ctypes.byref(fes.mesh.c_struct()), ctypes.byref(defo), _ptr(_lset_vec(integ.lset)), fes.order, integ.dt,
                                       integ.intorder(fes.order), ctypes.byref(poly), els.numel(), _ptr(els), _ptr(elvecs), None, None, _stream()))
        check(L.ngsx_gather_vec(fes.ndof, _ptr(dofptr), _ptr(pairs), _ptr(elvecs), 1.0, 1, _ptr(self.vec), _stream()))

    def _assemble_x(self, integ, k, gfd=None):
        """x-dof part (DiffOpX<RNEG/RPOS> as test operator, xfem/xfemdiffops.cpp:30-40): the element vectors of the IF elements
        (the only ones with x-dofs), gathered per x-dof over the entries whose sign matches the operator -- fixed order, no atomics"""
        L = _lib_()
        x, fes = self.space.x, self._h1
        nd = fes.nd
        if k not in self._xplans:
            keep = torch.ones(x.if_els.numel(), dtype=torch.bool, device=x.if_els.device)
            if integ.definedonelements is not None:
                keep = integ.definedonelements.to_bool()[x.if_els.long()]
            plans = []
            for blk, sign in ((1, NEG), (2, POS)):
                if integ.blocks[blk] == 0.0:
                    continue
                sel = ((x.xsign == sign) & keep[:, None]).reshape(-1)
                src = torch.nonzero(sel).reshape(-1)                                       # index into the flattened [nIF][nd] vectors
                dof = x.el2dofs.reshape(-1)[src].long()
                order = torch.argsort(dof * (x.if_els.numel() * nd) + src)                 # by x-dof, then by (element, local dof)
                dofptr = torch.zeros(x.ndof + 1, dtype=torch.int64, device=src.device)
                dofptr[1:] = torch.cumsum(torch.bincount(dof, minlength=x.ndof), 0)
                plans.append((integ.blocks[blk], dofptr, src[order].contiguous()))
            self._xplans[k] = (plans, torch.empty((x.if_els.numel(), nd), dtype=torch.float64, device=x.if_els.device))
        plans, elvecs = self._xplans[k]
        if x.if_els.numel() == 0:
            return
        poly = _poly_of(integ.poly)
        if gfd is not None:                                                # deformed mesh: shapes at the reference points, f at the mapped ones
            defo = fes.mesh.c_deformation(gfd)
            check(L.ngsx_elvec_h1_deformed(ctypes.byref(fes.mesh.c_struct()), ctypes.byref(defo), _ptr(_lset_vec(integ.lset)), fes.order, integ.dt,
                                           integ.intorder(fes.order), ctypes.byref(poly), x.if_els.numel(), _ptr(x.if_els), _ptr(elvecs), None, None,
                                           _stream()))
        else:
            check(L.ngsx_elvec_h1(ctypes.byref(fes.mesh.c_struct()), _ptr(_lset_vec(integ.lset)), fes.order, integ.dt, integ.intorder(fes.order),
                                  ctypes.byref(poly), x.if_els.numel(), _ptr(x.if_els), _ptr(elvecs), None, _stream()))
        xvec = self.vec[fes.ndof:]
        for scale, dofptr, pairs in plans:
            check(L.ngsx_gather_vec(x.ndof, _ptr(dofptr), _ptr(pairs), _ptr(elvecs), float(scale), 1, _ptr(xvec), _stream()))


# XFESpace and the sign-masked operators (xfem/python_xfem.cpp:321-448, python/__init__.py:22-105, 397-409)
from .xfes import XFESpace, CutRatioGF, neg, pos, neg_grad, pos_grad, extend, extend_grad  # noqa: E402,F401
from .spacetime import ScalarTimeFE, SpaceTimeFESpace, SpaceTimeInterpolateToP1, dt  # noqa: E402,F401
from .ngs import ReferenceTimeVariable  # noqa: E402,F401
