"""Minimal stand-ins for the NGSolve objects the xfem surface takes (Mesh, H1, GridFunction, BitArray,
CoefficientFunction algebra, BilinearForm / LinearForm) so that the hot path can be driven -- and tested -- with the
same call sequences as the reference's tutorials, without NGSolve.  Host-side logic only: every array lives in a
torch tensor on the CUDA device and every computation on the path goes through the C ABI (ngsxfem_b200._lib).

In an NGSolve embedding these classes are replaced by NGSolve's own (INTEGRATION.md shows the adaptor)."""
import math
from typing import Dict, List, Optional, Tuple

import torch

from . import _lib
from .meshes import SimplexMesh, TensorMesh

VOL, BND = 0, 1


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def require_cuda(t: torch.Tensor, what: str):
    if not t.is_cuda:
        raise _lib.NgsxError(-1, f"{what} must live on a CUDA device (there is no CPU path)")


# ------------------------------------------------------------------------------------------------------------------
class BitArray:
    """ngcore::BitArray layout: bit i = word i//32, bit i%32 (uint32 words stored as int32 tensor on the device)."""

    def __init__(self, size: int, words: Optional[torch.Tensor] = None, device=None):
        self.size = int(size)
        nw = (self.size + 31) // 32
        if words is None:
            words = torch.zeros(nw, dtype=torch.int32, device=device)
        assert words.dtype == torch.int32 and words.numel() == nw
        self.words = words

    def __len__(self):
        return self.size

    @property
    def device(self):
        return self.words.device

    def to_bool(self) -> torch.Tensor:
        idx = torch.arange(self.size, device=self.device)
        return ((self.words[idx >> 5] >> (idx & 31)) & 1).bool()

    @staticmethod
    def from_bool(mask: torch.Tensor) -> "BitArray":
        n = mask.numel()
        nw = (n + 31) // 32
        pad = torch.zeros(nw * 32, dtype=torch.int64, device=mask.device)
        pad[:n] = mask.long()
        w = (pad.reshape(nw, 32) << torch.arange(32, device=mask.device)).sum(dim=1)
        w = torch.where(w >= 2 ** 31, w - 2 ** 32, w).to(torch.int32)
        return BitArray(n, w)

    def NumSet(self) -> int:
        return int(self.to_bool().sum().item())

    def _tailmask(self):
        r = self.size & 31
        if r:
            m = (1 << r) - 1
            self.words[-1] = self.words[-1] & (m if m < 2 ** 31 else m - 2 ** 32)
        return self

    def Set(self):
        self.words.fill_(-1)
        return self._tailmask()

    def Clear(self):
        self.words.zero_()
        return self

    def __or__(self, o):
        return BitArray(self.size, self.words | o.words)

    def __and__(self, o):
        return BitArray(self.size, self.words & o.words)

    def __invert__(self):
        return BitArray(self.size, ~self.words)._tailmask()

    def __getitem__(self, i):
        return bool((int(self.words[i >> 5].item()) >> (i & 31)) & 1)

    def nonzero(self) -> torch.Tensor:
        return torch.nonzero(self.to_bool()).reshape(-1)


# ------------------------------------------------------------------------------------------------------------------
class QMesh(TensorMesh):
    """device-resident mesh of quadrilaterals / hexahedra (row f-5): cut quadrature only (CutInfo, Integrate, rule API)"""

    def __init__(self, m: TensorMesh, device="cuda"):
        super().__init__(m.dim, m.vert.to(device).contiguous(), m.el2v.to(device).contiguous())
        self._c = None
        self.deformation = None

    def c_struct(self) -> _lib.Mesh:
        if self._c is None:
            require_cuda(self.vert, "mesh")
            self._c = _lib.Mesh(self.dim, 0, self.ne, self.nv, self.vert.data_ptr(), self.el2v.data_ptr())
        return self._c

    def c_deformation(self, gf=None):
        return None

    def GetNE(self, vb=VOL):
        return self.ne


class Mesh(SimplexMesh):
    """device-resident simplex mesh + the ngsx_mesh struct of the C ABI (a TensorMesh gives a QMesh)"""

    def __new__(cls, m=None, device="cuda"):
        if isinstance(m, TensorMesh):
            return QMesh(m, device)
        return super().__new__(cls)

    def __init__(self, m: SimplexMesh, device="cuda"):
        super().__init__(m.dim, m.vert.to(device).contiguous(), m.el2v.to(device).contiguous())
        self._c = None
        self.deformation = None
        self._block = None

    def c_struct(self) -> _lib.Mesh:
        if self._c is None:
            require_cuda(self.vert, "mesh")
            self._c = _lib.Mesh(self.dim, 0, self.ne, self.nv, self.vert.data_ptr(), self.el2v.data_ptr())
        return self._c

    def GetNE(self, vb=VOL):
        return self.ne

    def element_block(self, lo: int, hi: int) -> "Mesh":
        """view on the elements [lo, hi) sharing the vertex array (the per-rank mesh block of element-wise reductions)"""
        m = Mesh.__new__(Mesh)
        SimplexMesh.__init__(m, self.dim, self.vert, self.el2v[lo:hi])
        m._c = None
        m.deformation, m._block = self.deformation, (lo, hi)
        return m

    def SetDeformation(self, gf: "GridFunction"):
        """MeshAccess::SetDeformation (NGS/comp/meshaccess.hpp): every ElementTransformation becomes the affine map plus the
        vector H1 field gf (H1(mesh, order = 1..2, dim = mesh.dim)); SURVEY section 8 row f-4"""
        if gf is not None:
            fes = gf.space
            if not isinstance(fes, H1) or fes.dim != self.dim or fes.order not in (1, 2):
                raise _lib.NgsxError(-2, "deformation must be a GridFunction on H1(mesh, order = 1 or 2, dim = mesh.dim); no CPU fallback")
        self.deformation = gf

    def UnsetDeformation(self):
        self.deformation = None

    def c_deformation(self, gf: "GridFunction" = None):
        """ngsx_deformation of gf (default: the mesh deformation), restricted to this element block; None if undeformed"""
        gf = gf if gf is not None else self.deformation
        if gf is None:
            return None
        fes = gf.space
        el2dof = fes.el2dof if self._block is None else fes.el2dof[self._block[0]:self._block[1]]
        coef = gf.vec.reshape(fes.ndof, self.dim)
        require_cuda(coef, "deformation")
        d = _lib.Deformation(fes.order, fes.nd, fes.ndof, el2dof.data_ptr(), coef.data_ptr())
        d._keep = (el2dof, coef)
        return d

    @property
    def nfacet(self):
        return self.facets()[1].shape[0]


class FESpace:
    pass


class H1(FESpace):
    """H1HighOrderFESpace stand-in (orders 1..2 on every entry point of the CUDA path, order 3 for bilinear / linear forms)."""

    def __init__(self, mesh: Mesh, order: int = 1, dgjumps: bool = False, dim: int = 1, **_):
        self.mesh, self.order, self.dgjumps, self.dim = mesh, order, dgjumps, dim          # dim > 1: interleaved vector field
        self.el2dof, self.ndof = mesh.h1_dofs(order)
        self.nd = self.el2dof.shape[1]

    def TrialFunction(self):
        return Expr.atom("u")

    def TestFunction(self):
        return Expr.atom("v")

    def TnT(self):
        return self.TrialFunction(), self.TestFunction()


class GridFunction:
    def __init__(self, fes: H1):
        self.space = fes
        shape = fes.ndof if getattr(fes, "dim", 1) == 1 else (fes.ndof, fes.dim)
        self.vec = torch.zeros(shape, dtype=torch.float64, device=fes.mesh.device)

    # a scalar H1 GridFunction of order 1-2 as a factor of the coefficient of a linear form:  gf * v,  poly * gf * v
    def __mul__(self, o):
        if isinstance(o, Expr):
            return o * self
        if isinstance(o, (int, float, Poly)):
            return _GfCoef(self, Poly.lift(o))
        return NotImplemented

    __rmul__ = __mul__


class _GfCoef:
    """poly * GridFunction: a coefficient waiting for its test function"""

    def __init__(self, gf, poly):
        self.gf, self.poly = gf, poly

    def __mul__(self, o):
        if isinstance(o, Expr):
            return (o * self.poly) * self.gf if not set(self.poly.terms) <= {(0, 0, 0)} else (o * self.poly.terms.get((0, 0, 0), 0.0)) * self.gf
        if isinstance(o, (int, float, Poly)):
            return _GfCoef(self.gf, self.poly * Poly.lift(o))
        return NotImplemented

    __rmul__ = __mul__


# ------------------------------------------------------------------------------------------------------------------
# coefficient-function algebra: just enough to write the forms of the tutorials and polynomial integrands
# ------------------------------------------------------------------------------------------------------------------
class Poly:
    """polynomial CoefficientFunction in x, y, z: {(ex,ey,ez): coef}"""

    def __init__(self, terms: Dict[Tuple[int, int, int], float]):
        self.terms = {k: float(v) for k, v in terms.items() if v != 0.0}

    @staticmethod
    def lift(o):
        if isinstance(o, Poly):
            return o
        if isinstance(o, (int, float)):
            return Poly({(0, 0, 0): float(o)})
        return None

    def __add__(self, o):
        o = Poly.lift(o)
        if o is None:
            return NotImplemented
        t = dict(self.terms)
        for k, v in o.terms.items():
            t[k] = t.get(k, 0.0) + v
        return Poly(t)

    __radd__ = __add__

    def __neg__(self):
        return Poly({k: -v for k, v in self.terms.items()})

    def __sub__(self, o):
        o = Poly.lift(o)
        return NotImplemented if o is None else self + (-o)

    def __rsub__(self, o):
        return (-self) + o

    def __mul__(self, o):
        p = Poly.lift(o)
        if p is None:
            return NotImplemented
        t = {}
        for (a, b, c), v in self.terms.items():
            for (d, e, f), w in p.terms.items():
                k = (a + d, b + e, c + f)
                t[k] = t.get(k, 0.0) + v * w
        return Poly(t)

    __rmul__ = __mul__

    def __pow__(self, n):
        r = Poly({(0, 0, 0): 1.0})
        for _ in range(int(n)):
            r = r * self
        return r

    def as_terms(self):
        return [(c, *k) for k, c in sorted(self.terms.items())] or [(0.0, 0, 0, 0)]

    def is_affine(self):
        return all(sum(k) <= 1 for k in self.terms)


x, y, z = Poly({(1, 0, 0): 1.0}), Poly({(0, 1, 0): 1.0}), Poly({(0, 0, 1): 1.0})


class TPoly:
    """polynomial in x, y, z and the reference time tref (ReferenceTimeVariable of the reference's space-time module):
    {(ex, ey, ez, et): coef}"""

    def __init__(self, terms):
        self.terms = {k: float(v) for k, v in terms.items() if v != 0.0}

    @staticmethod
    def lift(o):
        if isinstance(o, TPoly):
            return o
        p = Poly.lift(o)
        return None if p is None else TPoly({(*k, 0): v for k, v in p.terms.items()})

    def __add__(self, o):
        o = TPoly.lift(o)
        if o is None:
            return NotImplemented
        t = dict(self.terms)
        for k, v in o.terms.items():
            t[k] = t.get(k, 0.0) + v
        return TPoly(t)

    __radd__ = __add__

    def __neg__(self):
        return TPoly({k: -v for k, v in self.terms.items()})

    def __sub__(self, o):
        o = TPoly.lift(o)
        return NotImplemented if o is None else self + (-o)

    def __rsub__(self, o):
        return (-self) + o

    def __mul__(self, o):
        o = TPoly.lift(o)
        if o is None:
            return NotImplemented
        t = {}
        for k1, v in self.terms.items():
            for k2, w in o.terms.items():
                k = tuple(a + b for a, b in zip(k1, k2))
                t[k] = t.get(k, 0.0) + v * w
        return TPoly(t)

    __rmul__ = __mul__

    def __pow__(self, n):
        r = TPoly({(0, 0, 0, 0): 1.0})
        for _ in range(int(n)):
            r = r * self
        return r

    def as_terms(self):
        return [(c, *k) for k, c in sorted(self.terms.items())] or [(0.0, 0, 0, 0, 0)]


def ReferenceTimeVariable():
    """tref of the reference (spacetime/python_spacetime.cpp): the reference time in [0, 1] of a space-time quadrature point"""
    return TPoly({(0, 0, 0, 1): 1.0})


def CoefficientFunction(v):
    return Poly.lift(v)


class AnalyticLevelset:
    """level-set CoefficientFunction evaluated by the ngsx_interpolate_p1_cf kernel"""

    def __init__(self, kind: int, params):
        self.kind = kind
        self.params = list(params) + [0.0] * (8 - len(params))


def SphereLevelset(center, radius):
    c = list(center) + [0.0] * (3 - len(center))
    return AnalyticLevelset(_lib.LSET_SPHERE, [*c, radius])


def RingLevelset(center, r1, r2):
    """IfPos(r-rc, r-rc-rr, rc-r-rr) of py_tutorials/fictdom_ghostpen.py:28-33"""
    c = list(center) + [0.0] * (3 - len(center))
    return AnalyticLevelset(_lib.LSET_RING, [*c, (r1 + r2) / 2.0, (r2 - r1) / 2.0])


def CylinderZLevelset(center, radius):
    return AnalyticLevelset(_lib.LSET_CYLZ, [center[0], center[1], 0.0, radius])


class Expr:
    """sum of terms  coef * h^hpow * (product of atoms).  Atoms: u, v (scalars), gu, gv, n (vectors),
    and the contracted scalars gugv, gun, gvn; ju, jv = u - u.Other(), v - v.Other()."""

    def __init__(self, terms: List[Tuple[float, int, Tuple[str, ...]]], poly: Optional[Poly] = None, gf=None):
        self.terms = terms
        self.poly = poly          # optional polynomial coefficient multiplying the whole expression (f * v)
        self.gf = gf              # optional GridFunction factor of that coefficient (scalar H1, order 1-2)

    @staticmethod
    def atom(name):
        return Expr([(1.0, 0, (name,))])

    def Other(self):
        (c, p, a), = self.terms
        assert a in (("u",), ("v",)) and c == 1.0 and p == 0
        return Expr([(1.0, 0, (a[0] + "o",))])

    @staticmethod
    def lift(o):
        if isinstance(o, Expr):
            return o
        if isinstance(o, (int, float)):
            return Expr([(float(o), 0, ())])
        if isinstance(o, Poly) and set(o.terms) <= {(0, 0, 0)}:
            return Expr([(o.terms.get((0, 0, 0), 0.0), 0, ())])
        return None

    def __add__(self, o):
        o = Expr.lift(o)
        return NotImplemented if o is None else Expr(self.terms + o.terms)

    __radd__ = __add__

    def __neg__(self):
        return Expr([(-c, p, a) for c, p, a in self.terms])

    def __sub__(self, o):
        o = Expr.lift(o)
        return NotImplemented if o is None else self + (-o)

    def __rsub__(self, o):
        return (-self) + o

    # vector-valued atoms: gradients of the trial / test function on the dofs of Vh (gu, gv), the sign-masked gradients on the
    # x-dofs (gxun = neg_grad(u_x), gxup = pos_grad(u_x), gxvn, gxvp; xfem/xfemdiffops.cpp:44-60) and the level-set normal n
    _VEC = {"gu", "gv", "n", "gxun", "gxup", "gxvn", "gxvp"}
    _CONTRACT = {frozenset(("gu", "gv")): "gugv", frozenset(("gu", "n")): "gun", frozenset(("gv", "n")): "gvn"}

    @staticmethod
    def _contract(vec):
        if list(vec) == ["n", "n"]:
            return None                                # |n| = 1
        key = frozenset(vec)
        if key in Expr._CONTRACT:
            return Expr._CONTRACT[key]
        a, b = sorted(vec)
        trial, test = {"gu", "gxun", "gxup"}, {"gv", "gxvn", "gxvp"}
        ok = (a in trial and b in test) or (a in test and b in trial) or ((a == "n") != (b == "n"))
        if not ok or len(key) != 2:
            raise _lib.NgsxError(-2, f"unsupported product {vec}")
        return a + "." + b

    def __mul__(self, o):
        if isinstance(o, GridFunction):
            if self.gf is not None or getattr(o.space, "dim", 1) != 1 or not isinstance(o.space, H1) or o.space.order not in (1, 2):
                raise _lib.NgsxError(-2, "one scalar H1 GridFunction of order 1-2 may multiply a linear form f*v")
            return Expr(list(self.terms), self.poly, o)
        if isinstance(o, _GfCoef):
            return o * self
        if isinstance(o, Poly) and not set(o.terms) <= {(0, 0, 0)}:
            return Expr(list(self.terms), o if self.poly is None else self.poly * o, self.gf)
        o = Expr.lift(o)
        if o is None:
            return NotImplemented
        if self.gf is not None or o.gf is not None:
            if o.terms and all(a == () and p == 0 for _, p, a in o.terms) and o.gf is None:      # scalar * (gf * v)
                c = sum(cc for cc, _, _ in o.terms)
                return Expr([(c * c1, p1, a1) for c1, p1, a1 in self.terms], self.poly, self.gf)
            raise _lib.NgsxError(-2, "a GridFunction coefficient may only multiply a finished linear form f*v")
        if self.poly is not None or o.poly is not None:
            raise _lib.NgsxError(-2, "a polynomial coefficient may only multiply a finished linear form f*v")
        out = []
        for c1, p1, a1 in self.terms:
            for c2, p2, a2 in o.terms:
                atoms = list(a1 + a2)
                vec = [a for a in atoms if a in Expr._VEC]
                if len(vec) == 2:                      # NGSolve: vector * vector = inner product
                    c = Expr._contract(vec)
                    atoms = [a for a in atoms if a not in Expr._VEC] + ([c] if c is not None else [])
                elif len(vec) > 2:
                    raise _lib.NgsxError(-2, "unsupported vector product")
                out.append((c1 * c2, p1 + p2, tuple(sorted(atoms))))
        return Expr(out)

    __rmul__ = __mul__

    def __truediv__(self, o):
        o = Expr.lift(o)
        if o is None or len(o.terms) != 1 or o.terms[0][2] != ():
            raise _lib.NgsxError(-2, "division only by scalars / powers of h")
        c, p, _ = o.terms[0]
        return Expr([(c1 / c, p1 - p, a1) for c1, p1, a1 in self.terms])

    def __rtruediv__(self, o):
        return Expr.lift(o) / self

    def collect(self) -> Dict[Tuple[str, ...], Dict[int, float]]:
        """expand ju/jv/uo/vo and sum like terms: {atoms: {hpow: coef}}"""
        res: Dict[Tuple[str, ...], Dict[int, float]] = {}
        for c, p, a in self.terms:
            d = res.setdefault(a, {})
            d[p] = d.get(p, 0.0) + c
        return {a: {p: c for p, c in d.items() if c != 0.0} for a, d in res.items() if any(v != 0.0 for v in d.values())}


class _MeshSize(Expr):
    def __init__(self):
        super().__init__([(1.0, 1, ())])


class specialcf:
    mesh_size = _MeshSize()


def grad(e: Expr) -> Expr:
    (c, p, a), = e.terms
    if a == ("u",):
        return Expr([(c, p, ("gu",))])
    if a == ("v",):
        return Expr([(c, p, ("gv",))])
    raise _lib.NgsxError(-2, "grad() of this expression is outside the CUDA path")


def LevelsetNormal(_lsetp1=None) -> Expr:
    """n = grad(lsetp1)/|grad(lsetp1)| (py_tutorials/fictdom_ghostpen.py:71); elementwise constant for a P1 level set"""
    return Expr.atom("n")


class CSRMatrix:
    """SparseMatrix<double> stand-in: rowptr int64 [ndof+1], colidx int32 [nnz], vals float64 [nnz] on the device"""

    def __init__(self, ndof, rowptr, colidx, vals):
        self.ndof, self.rowptr, self.colidx, self.vals = ndof, rowptr, colidx, vals

    @property
    def nnz(self):
        return self.colidx.numel()

    def to_scipy(self):
        import scipy.sparse as sp
        return sp.csr_matrix((self.vals.cpu().numpy(), self.colidx.cpu().numpy(), self.rowptr.cpu().numpy()),
                             shape=(self.ndof, self.ndof))
