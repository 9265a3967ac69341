"""XFESpace and the sign-masked operators of the reference on the CUDA path.

Mirrors xfem/xFESpace.cpp:244-366 (T_XFESpace::Update: x-dofs = base dofs of the IF elements, domofdof = INVERT of the node's
domain), xfem/python_xfem.cpp:321-448 (XFESpace(basefes, cutinfo | lset), GetCutInfo, BaseDofOfXDof, GetDomainOfDof,
GetDomainNrs), python/__init__.py:22-105 (neg, pos, neg_grad, pos_grad), :397-409 (CutRatioGF) and NGSolve's compound space
FESpace([Vh, Vhx]) as far as py_tutorials/nxfem.py:62-107 uses it.  Host-side bookkeeping only (torch tensors on the device);
the element matrices come from ngsx_xfem_elmat_h1, the scatter from ngsx_assemble_csr_rows_mats."""
import ctypes
from typing import List, Optional

import torch

from . import _lib
from ._lib import NEG, POS, IF, NgsxError, check
from .ngs import Expr, FESpace as _FESpaceBase, GridFunction, H1, _ptr, _stream


class XProxy:
    """trial / test function of an XFESpace component: only neg / pos / neg_grad / pos_grad of it are meaningful"""

    def __init__(self, kind: str):
        self.kind = kind            # "u" or "v"


def neg(f: XProxy) -> Expr:
    """python/__init__.py:51-63 -> DiffOpX<D, RNEG> (xfem/xfemdiffops.cpp:30-40)"""
    return Expr.atom("x" + _kind(f) + "n")


def pos(f: XProxy) -> Expr:
    """python/__init__.py:37-49 -> DiffOpX<D, RPOS>"""
    return Expr.atom("x" + _kind(f) + "p")


def neg_grad(f: XProxy) -> Expr:
    """python/__init__.py:93-105 -> DiffOpX<D, RNEG_GRAD> (xfem/xfemdiffops.cpp:49-60)"""
    return Expr.atom("gx" + _kind(f) + "n")


def pos_grad(f: XProxy) -> Expr:
    """python/__init__.py:79-91 -> DiffOpX<D, RPOS_GRAD>"""
    return Expr.atom("gx" + _kind(f) + "p")


def extend(f):
    raise NgsxError(-2, "extend / extend_grad (DiffOpX EXTEND) are outside the CUDA path")


extend_grad = extend


def _kind(f):
    if not isinstance(f, XProxy):
        raise NgsxError(-1, "neg / pos / neg_grad / pos_grad take the trial or test function of an XFESpace")
    return f.kind


def CutRatioGF(cutinfo) -> Expr:
    """python/__init__.py:397-409: elementwise constant cut ratio as a coefficient (atom k0; 1 - CutRatioGF expands symbolically)"""
    e = Expr.atom("k0")
    e.cutinfo = cutinfo
    return e


class XFESpace(_FESpaceBase):
    """xfem/xFESpace.cpp: enrichment space on the IF elements of a CutInfo"""

    def __init__(self, basefes: H1, cutinfo=None, lset=None, flags=None, heapsize: int = 1000000):
        from .xfem import CutInfo
        if not isinstance(basefes, H1):
            raise NgsxError(-2, "XFESpace: only scalar H1 base spaces are on the CUDA path")
        if isinstance(cutinfo, GridFunction) and lset is None:            # python_xfem.cpp:331-336: a level set in the cutinfo slot
            cutinfo, lset = None, cutinfo
        if cutinfo is None and lset is None:
            raise NgsxError(-1, "levelset and cutinfo are invalid")
        self.basefes, self.mesh, self.order = basefes, basefes.mesh, basefes.order
        self.private_cutinfo = cutinfo is None
        self.lset = lset
        self.cutinfo = cutinfo if cutinfo is not None else CutInfo(self.mesh, lset)
        self.Update()

    def Update(self):
        from .xfem import GetDofsOfElements
        ci, base = self.cutinfo, self.basefes
        if self.private_cutinfo:
            ci.Update(self.lset)                                           # xFESpace.cpp:248-253
        ci._node_marks()
        dev = self.mesh.device
        active = GetDofsOfElements(base, ci.GetElementsOfType(IF)).to_bool()          # xFESpace.cpp:276-295
        self.ndof = int(active.sum().item())
        csum = torch.cumsum(active.to(torch.int64), 0) - 1
        self.basedof2xdof = torch.where(active, csum, torch.full_like(csum, -1)).to(torch.int32)      # :303-311
        self.xdof2basedof = torch.nonzero(active).reshape(-1).to(torch.int32)                          # :313-321
        # domain of dof (:346-366): NEG unless the dof's node touches an uncut element, then INVERT(domain of that node)
        node_dom = [ci.dom_of_node["vertex"]]
        if base.order >= 2:
            node_dom.append(ci.dom_of_node["edge"].repeat_interleave(base.order - 1))
        node_dom = torch.cat(node_dom)[: base.ndof]
        if node_dom.numel() != base.ndof:
            raise NgsxError(-2, "XFESpace: base space with face / cell dofs is outside the CUDA path")
        d = node_dom[self.xdof2basedof.long()]
        self.domofdof = torch.where(d == IF, torch.full_like(d, NEG), 1 - d).to(torch.int8)
        self.if_els = ci.GetElementsOfType(IF).nonzero().to(torch.int32).contiguous()
        self.el2dofs = self.basedof2xdof[base.el2dof[self.if_els.long()].long()].contiguous()          # [nIF][nd] (:323-331)
        self.xsign = self.domofdof[self.el2dofs.long()].contiguous()                                   # GetDomainNrs per IF element
        self._elpos = None
        self.version = getattr(self, "version", 0) + 1      # forms compare it and rebuild what they derived from the cut state
        return self

    def GetCutInfo(self):
        return self.cutinfo

    def BaseDofOfXDof(self, i: int) -> int:
        return int(self.xdof2basedof[i].item())

    def GetDomainOfDof(self, i: int) -> int:
        return int(self.domofdof[i].item())

    def _pos_of(self, elnr: int) -> int:
        if self._elpos is None:
            self._elpos = {int(e): k for k, e in enumerate(self.if_els.tolist())}
        return self._elpos.get(int(elnr), -1)

    def GetDofNrs(self, elnr: int) -> List[int]:
        k = self._pos_of(elnr)
        return [] if k < 0 else self.el2dofs[k].tolist()                  # xFESpace.cpp:15-28

    def GetDomainNrs(self, elnr: int) -> List[int]:
        k = self._pos_of(elnr)
        return [] if k < 0 else self.xsign[k].tolist()                    # xFESpace.cpp:30-53

    def TrialFunction(self):
        return XProxy("u")

    def TestFunction(self):
        return XProxy("v")

    def TnT(self):
        return self.TrialFunction(), self.TestFunction()


class CompoundFESpace(_FESpaceBase):
    """FESpace([Vh, Vhx]): dofs of Vh first, then the x-dofs (NGSolve CompoundFESpace numbering)"""
    dgjumps = False

    def __init__(self, spaces):
        if len(spaces) != 2 or not isinstance(spaces[0], H1) or not isinstance(spaces[1], XFESpace) or spaces[1].basefes is not spaces[0]:
            raise NgsxError(-2, "compound spaces other than [Vh, XFESpace(Vh)] are outside the CUDA path")
        self.components = list(spaces)
        self.std, self.x = spaces
        self.mesh, self.order = self.std.mesh, self.std.order

    @property
    def ndof(self):
        return self.std.ndof + self.x.ndof

    def compound_table(self) -> torch.Tensor:
        """[nIF][2 nd]: dofs of Vh, then offset x-dofs, of every IF element (CompoundFESpace::GetDofNrs)"""
        els = self.x.if_els.long()
        return torch.cat([self.std.el2dof[els].to(torch.int32), self.x.el2dofs + self.std.ndof], dim=1).contiguous()

    def TrialFunction(self):
        return self.std.TrialFunction(), self.x.TrialFunction()

    def TestFunction(self):
        return self.std.TestFunction(), self.x.TestFunction()

    def TnT(self):
        return self.TrialFunction(), self.TestFunction()


def FESpace(spaces, **_):
    """ngsolve.FESpace([...]) as used by py_tutorials/nxfem.py:63"""
    return CompoundFESpace(spaces)


# ------------------------------------------------------------------------------------------------------------------
# form -> product terms of the C ABI (ngsx_xterm)
# ------------------------------------------------------------------------------------------------------------------
_TRIAL = {"u": (0, "val"), "xun": (1, "val"), "xup": (2, "val"), "gun": (0, "dn"), "gxun.n": (1, "dn"), "gxup.n": (2, "dn")}
_TEST = {"v": (0, "val"), "xvn": (1, "val"), "xvp": (2, "val"), "gvn": (0, "dn"), "gxvn.n": (1, "dn"), "gxvp.n": (2, "dn")}
_GBLK = {"gu": 0, "gxun": 1, "gxup": 2, "gv": 0, "gxvn": 1, "gxvp": 2}
_PAIR = {("val", "val"): 0, ("val", "dn"): 1, ("dn", "val"): 2, ("dn", "dn"): 3}     # (test, trial)


def xterms_of(form: Expr):
    """expand a bilinear form on [Vh, Vhx] into (coef, hpow, kappa, pair, blk_u, blk_v) tuples and the differential orders
    of its trial / test proxies (xfem/symboliccutbfi.cpp:97-101; DiffOpX::DIFFORDER, xfem/xfemdiffops.hpp:24)"""
    out = []
    do_u, do_v = 99, 99
    for atoms, d in form.collect().items():
        atoms = list(atoms)
        nk = atoms.count("k0")
        if nk > 1:
            raise NgsxError(-2, "products of cut ratios are outside the CUDA path")
        atoms = [a for a in atoms if a != "k0"]
        if len(atoms) == 1:
            a = atoms[0]
            parts = ["gu", "gv"] if a == "gugv" else a.split(".")
            tu = [x for x in parts if x in ("gu", "gxun", "gxup")]
            tv = [x for x in parts if x in ("gv", "gxvn", "gxvp")]
            if len(parts) != 2 or len(tu) != 1 or len(tv) != 1:
                raise NgsxError(-2, f"form term {atoms} is outside the CUDA path")
            pair, bu, bv = 4, _GBLK[tu[0]], _GBLK[tv[0]]
            do_u, do_v = min(do_u, 1), min(do_v, 1)
        elif len(atoms) == 2:
            tu = [x for x in atoms if x in _TRIAL]
            tv = [x for x in atoms if x in _TEST]
            if len(tu) != 1 or len(tv) != 1:
                raise NgsxError(-2, f"form term {atoms} is outside the CUDA path")
            (bu, ku), (bv, kv) = _TRIAL[tu[0]], _TEST[tv[0]]
            pair = _PAIR[(kv, ku)]
            do_u, do_v = min(do_u, 0 if ku == "val" else 1), min(do_v, 0 if kv == "val" else 1)
        else:
            raise NgsxError(-2, f"form term {atoms} is outside the CUDA path (one trial and one test factor per term)")
        for p, c in d.items():
            out.append((c, p, 1 if nk else 0, pair, bu, bv))
    return out, (do_u, do_v)


def make_xterms(terms):
    arr = (_lib.XTerm * max(1, len(terms)))()
    for t, (c, p, k, pair, bu, bv) in enumerate(terms):
        arr[t].coef, arr[t].hpow, arr[t].kappa, arr[t].pair, arr[t].blk_u, arr[t].blk_v = float(c), int(p), int(k), int(pair), int(bu), int(bv)
    return arr


def has_x_atoms(form: Expr) -> bool:
    names = ("xun", "xup", "xvn", "xvp", "gxun", "gxup", "gxvn", "gxvp", "k0")
    return any(any(any(n in a.split(".") for n in names) for a in atoms) for _, _, atoms in form.terms)
