"""Space-time level sets and cut quadrature on the CUDA path: stand-ins for ScalarTimeFE,
SpaceTimeFESpace and SpaceTimeInterpolateToP1 (spacetime/python_spacetime.cpp, spacetime/SpaceTimeFESpace.cpp) as far as the
space-time cut rule needs them -- a level set that is P1 in space and nodal in the reference time.  The quadrature itself is
ngsx_st_rule_count / _fill / ngsx_st_integrate (cutint/spacetimecutrule.cpp:98-196)."""
import ctypes

import torch

from . import _lib
from ._lib import NgsxError, check
from .ngs import Expr, FESpace, GridFunction, H1, Poly, TPoly, _ptr, _stream

_S5, _S37, _S7 = 5.0 ** 0.5, (3.0 / 7.0) ** 0.5, 7.0 ** 0.5
_NODES = {0: [0.0], 1: [0.0, 1.0], 2: [0.0, 0.5, 1.0],          # NodalTimeFE::CalcInterpolationPoints, spacetime/SpaceTimeFE.cpp:150-173
          3: [0.0, 0.5 * (1.0 - 1.0 / _S5), 0.5 * (1.0 + 1.0 / _S5), 1.0],
          4: [0.0, 0.5 * (1.0 - _S37), 0.5, 0.5 * (1.0 + _S37), 1.0],
          5: [0.0, 0.5 * (1.0 - (1.0 / 3.0 + 2.0 * _S7 / 21.0) ** 0.5), 0.5 * (1.0 - (1.0 / 3.0 - 2.0 * _S7 / 21.0) ** 0.5),
              0.5 * (1.0 + (1.0 / 3.0 - 2.0 * _S7 / 21.0) ** 0.5), 0.5 * (1.0 + (1.0 / 3.0 + 2.0 * _S7 / 21.0) ** 0.5), 1.0]}


class ScalarTimeFE:
    """spacetime/python_spacetime.cpp: nodal time element of the given order (0..5, the orders NodalTimeFE has nodes for)"""

    def __init__(self, order: int = 1, skip_first_node: bool = False, only_first_node: bool = False):
        if skip_first_node or only_first_node or order not in _NODES:
            raise NgsxError(-2, "ScalarTimeFE: orders 0..5 without node skipping are on the CUDA path")
        self.order = order
        self.nodes = _NODES[order]
        self.ndof = order + 1


class SpaceTimeFESpace(FESpace):
    """tensor-product space: dof = k * ndof_space + space dof (spacetime/SpaceTimeFESpace.cpp:109-122)"""

    def __init__(self, space: H1, tfe: ScalarTimeFE, **_):
        if not isinstance(space, H1) or space.order not in (1, 2):
            raise NgsxError(-2, "SpaceTimeFESpace: H1 of order 1..2 in space is on the CUDA path (the level set must be P1 in space)")
        self.space, self.tfe, self.mesh, self.order = space, tfe, space.mesh, space.order
        self.ndof = space.ndof * tfe.ndof
        self.nd = space.nd * tfe.ndof
        self._el2dof = None

    @property
    def el2dof(self):
        """[ne][ntd * nd_space]: dof (j, i) of SpaceTimeFE = j * ndof_space + space dof i (SpaceTimeFESpace::GetDofNrs)"""
        if self._el2dof is None:
            s = self.space.el2dof.long()
            self._el2dof = torch.cat([s + j * self.space.ndof for j in range(self.tfe.ndof)], dim=1).to(torch.int32).contiguous()
        return self._el2dof

    def TimeFE_nodes(self):
        return list(self.tfe.nodes)

    def TrialFunction(self):
        return Expr.atom("u")

    def TestFunction(self):
        return Expr.atom("v")

    def TnT(self):
        return self.TrialFunction(), self.TestFunction()


def SpaceTimeInterpolateToP1(spacetime_cf, time=None, spacetime_gf: GridFunction = None, eps_perturbation: float = 0.0):
    """nodal in time, P1 in space (SpaceTimeFESpace::InterpolateToP1): spacetime_cf is a polynomial in x, y, z, tref (TPoly) or a
    callable f(x, y[, z], t) on torch tensors; `time` (the reference's tref argument) is accepted for signature compatibility"""
    fes = spacetime_gf.space
    if not isinstance(fes, SpaceTimeFESpace):
        raise NgsxError(-1, "not a spacetime gridfunction")
    mesh = fes.mesh
    nv = mesh.nv
    cols = [mesh.vert[:, d] for d in range(mesh.dim)]
    for k, t in enumerate(fes.tfe.nodes):
        if callable(spacetime_cf):
            vals = spacetime_cf(*cols, t)
        else:
            p = TPoly.lift(spacetime_cf)
            if p is None:
                raise NgsxError(-2, "SpaceTimeInterpolateToP1: unsupported coefficient")
            c3 = cols + [torch.zeros_like(cols[0])] * (3 - len(cols))
            vals = torch.zeros_like(cols[0])
            for (a, b, c, d), coef in p.terms.items():
                vals = vals + coef * c3[0] ** a * c3[1] ** b * c3[2] ** c * (t ** d)
        vals = vals.to(torch.float64)
        if eps_perturbation > 0:
            vals = torch.where(vals.abs() < eps_perturbation, torch.full_like(vals, eps_perturbation), vals)
        spacetime_gf.vec[k * nv:(k + 1) * nv] = vals


def lset_st(levelset):
    """(device vector, number of time dofs) of a level-set GridFunction for the space-time rule; a plain P1 GridFunction is
    constant in time (the reference warns and does the same, cutint/xintegration.cpp:44-55)"""
    if isinstance(levelset, GridFunction) and isinstance(levelset.space, SpaceTimeFESpace):
        if levelset.space.order != 1:
            raise NgsxError(-2, "the space-time level set must be P1 in space")
        return levelset.vec, levelset.space.tfe.ndof
    if isinstance(levelset, GridFunction) and levelset.space.order == 1:
        return levelset.vec, 1
    raise NgsxError(-2, "Space-time requires the levelset as a GridFunction!")


def st_poly(cf):
    """polynomial integrand in (x, y, z, tref) -> (ngsx_poly, time exponents)"""
    p = TPoly.lift(1.0 if cf is None else cf)
    if p is None:
        raise NgsxError(-2, "only polynomial coefficient functions in x, y, z, tref are on the CUDA path")
    terms = p.as_terms()
    poly = _lib.make_poly([(c, a, b, d) for c, a, b, d, _ in terms])
    texp = (ctypes.c_int32 * len(terms))(*[int(t[4]) for t in terms])
    return poly, texp


def integrate(levelset, mesh, cf, order, domain_type, time_order, sync=True):
    vec, ntd = lset_st(levelset)
    poly, texp = st_poly(cf)
    out = torch.zeros(1, dtype=torch.float64, device=mesh.device)
    check(_lib.load().ngsx_st_integrate(ctypes.byref(mesh.c_struct()), _ptr(vec), ntd, int(domain_type), order, time_order,
                                        ctypes.byref(poly), texp, _ptr(out), _stream()))
    return float(out.item()) if sync else out


def cut_rule(levelset, mesh, domain_type, order_space, order_time):
    """the batched space-time rule: (npts [ne] int32, offs [ne+1] int64, points [n][3], tref [n], weights [n])"""
    L = _lib.load()
    vec, ntd = lset_st(levelset)
    dev = mesh.device
    mc = ctypes.byref(mesh.c_struct())
    npts = torch.zeros(mesh.ne, dtype=torch.int32, device=dev)
    check(L.ngsx_st_rule_count(mc, _ptr(vec), ntd, int(domain_type), order_space, order_time, _ptr(npts), _stream()))
    offs = torch.zeros(mesh.ne + 1, dtype=torch.int64, device=dev)
    check(L.ngsx_exclusive_scan(mesh.ne, _ptr(npts), _ptr(offs), _stream()))
    n = int(offs[-1].item())
    pts = torch.zeros((max(n, 1), 3), dtype=torch.float64, device=dev)
    tq = torch.zeros(max(n, 1), dtype=torch.float64, device=dev)
    w = torch.zeros(max(n, 1), dtype=torch.float64, device=dev)
    check(L.ngsx_st_rule_fill(mc, _ptr(vec), ntd, int(domain_type), order_space, order_time, _ptr(offs), _ptr(pts), _ptr(tq), _ptr(w), _stream()))
    return npts, offs, pts[:n], tq[:n], w[:n]


def classify(levelset, mesh, time_order):
    """CutInformation::Update with time_order >= 0 (xfem/cutinfo.cpp:79-111): (dom [ne] int8, cut ratio [ne])"""
    vec, ntd = lset_st(levelset)
    dom = torch.zeros(mesh.ne, dtype=torch.int8, device=mesh.device)
    ratio = torch.zeros(mesh.ne, dtype=torch.float64, device=mesh.device)
    check(_lib.load().ngsx_st_classify(ctypes.byref(mesh.c_struct()), _ptr(vec), ntd, int(time_order), _ptr(dom), _ptr(ratio), _stream()))
    return dom, ratio


def elmats(fes: SpaceTimeFESpace, levelset, domain_type, c_mass=0.0, c_stiff=0.0, c_dt=0.0, order_space=None, order_time=None, els=None,
           out=None):
    """space-time element matrices of  c_mass u v + c_stiff grad(u) grad(v) + c_dt dt(u) v  on SpaceTimeFE (spacetime/SpaceTimeFE.cpp:29-113)
    with the space-time cut rule: [n][N][N], N = fes.nd, rows = test functions.  Default orders as SymbolicCutBFI on a space-time
    space would pick them: 2 * order in space and in time."""
    mesh = fes.mesh
    vec, ntd = lset_st(levelset)
    os_ = 2 * fes.space.order if order_space is None else order_space
    ot = 2 * fes.tfe.order if order_time is None else order_time
    n = mesh.ne if els is None else els.numel()
    if out is None:
        out = torch.empty((n, fes.nd, fes.nd), dtype=torch.float64, device=mesh.device)
    check(_lib.load().ngsx_st_elmat_h1(ctypes.byref(mesh.c_struct()), _ptr(vec), ntd, fes.space.order, fes.tfe.ndof, int(domain_type), os_, ot,
                                       float(c_mass), float(c_stiff), float(c_dt), n, _ptr(els), _ptr(out), _stream()))
    return out


class SpaceTimeOperator:
    """assembled space-time operator of one cut integrator on SpaceTimeFESpace: CSR pattern from the space-time dof table
    (GetGraph), owner-computes row plan, element blocks from ngsx_st_elmat_h1, gather without atomics (ngsx_assemble_csr_rows_mats).
    Pattern and plan are built once; Assemble() re-evaluates the blocks for the current level set."""

    def __init__(self, fes: SpaceTimeFESpace, levelset, domain_type, c_mass=0.0, c_stiff=0.0, c_dt=0.0, order_space=None, order_time=None,
                 share: "SpaceTimeOperator" = None):
        from .ngs import CSRMatrix
        L = _lib.load()
        self.fes, self.levelset, self.dt = fes, levelset, int(domain_type)
        self.coef = (c_mass, c_stiff, c_dt)
        self.orders = (order_space, order_time)
        tab = fes.el2dof
        dev = fes.mesh.device
        self._owner = share is None
        if share is not None:                                # further integrators of one form: same pattern, plan and value array
            self.mat, self.plan = share.mat, share.plan
            self.blocks = torch.empty((tab.shape[0], fes.nd, fes.nd), dtype=torch.float64, device=dev)
            return
        g = ctypes.c_void_p()
        check(L.ngsx_csr_graph_create(fes.ndof, ctypes.byref(g)))
        try:
            check(L.ngsx_csr_graph_add_table(g, tab.shape[0], tab.shape[1], _ptr(tab), _stream()))
            nnz = ctypes.c_int64(0)
            check(L.ngsx_csr_graph_finalize(g, ctypes.byref(nnz), _stream()))
            rowptr = torch.empty(fes.ndof + 1, dtype=torch.int64, device=dev)
            colidx = torch.empty(nnz.value, dtype=torch.int32, device=dev)
            check(L.ngsx_csr_graph_export(g, _ptr(rowptr), _ptr(colidx), _stream()))
            torch.cuda.current_stream().synchronize()
        finally:
            L.ngsx_csr_graph_destroy(g)
        self.mat = CSRMatrix(fes.ndof, rowptr, colidx, torch.zeros(nnz.value, dtype=torch.float64, device=dev))
        h = ctypes.c_void_p()
        check(L.ngsx_rowplan_create(tab.shape[0], tab.shape[1], _ptr(tab), fes.ndof, _ptr(rowptr), _ptr(colidx), 0, fes.ndof, ctypes.byref(h),
                                    _stream()))
        self.plan = h
        self.blocks = torch.empty((tab.shape[0], fes.nd, fes.nd), dtype=torch.float64, device=dev)

    def Assemble(self, accumulate: bool = False):
        elmats(self.fes, self.levelset, self.dt, *self.coef, order_space=self.orders[0], order_time=self.orders[1], out=self.blocks)
        check(_lib.load().ngsx_assemble_csr_rows_mats(self.plan, _ptr(self.blocks), 1 if accumulate else 0, _ptr(self.mat.vals), _stream()))
        return self

    def __del__(self):
        try:
            if self.plan is not None and self._owner:
                _lib.load().ngsx_rowplan_destroy(self.plan)
                self.plan = None
        except Exception:
            pass


def dt(e: Expr) -> Expr:
    """spacetime/python_spacetime.cpp:194-249: derivative of a space-time trial function with respect to the REFERENCE time
    (SpaceTimeFE::CalcDtShape).  The CUDA path serves dt(u) * v."""
    (c, p, a), = e.terms
    if a == ("u",):
        return Expr([(c, p, ("dtu",))])
    raise NgsxError(-2, "dt() of this expression is outside the CUDA path (served: dt(u) on the trial function)")


class SymbolicSpaceTimeBFI:
    """SymbolicBFI(levelset_domain=..., form=..., time_order=k) on a SpaceTimeFESpace: the space-time cut rule of order k in time
    (cutint/xintegration.cpp:38-58) with the forms  c_m u v + c_s grad(u) grad(v) + c_t dt(u) v"""

    def __init__(self, lset, domain_type, form: Expr, time_order: int, force_intorder: int = -1):
        self.lset, self.dt, self.time_order, self.force_intorder = lset, int(domain_type), int(time_order), force_intorder
        self.coef = [0.0, 0.0, 0.0]
        names = {("u", "v"): 0, ("gugv",): 1, ("dtu", "v"): 2}
        if not isinstance(form, Expr) or form.poly is not None or getattr(form, "gf", None) is not None:
            raise NgsxError(-2, "space-time forms: c_m u v + c_s grad(u) grad(v) + c_t dt(u) v with constant coefficients; no CPU fallback")
        for atoms, d in form.collect().items():
            if atoms not in names or set(d) != {0}:
                raise NgsxError(-2, f"space-time form term {atoms} is outside the CUDA path (served: u*v, grad(u)*grad(v), dt(u)*v)")
            self.coef[names[atoms]] += d[0]


class SpaceTimeBilinearForm:
    """BilinearForm on a SpaceTimeFESpace: every integrator's blocks (ngsx_st_elmat_h1) gathered into one CSR matrix over the space-time
    dofs (pattern and row plan built once, owner-computes gather, no atomics)"""

    def __init__(self, space: SpaceTimeFESpace):
        self.space = space
        self.integrators = []
        self._ops = []
        self.mat = None

    def __iadd__(self, integ):
        if not isinstance(integ, SymbolicSpaceTimeBFI):
            raise NgsxError(-2, "a BilinearForm on a SpaceTimeFESpace takes SymbolicBFI(levelset_domain=..., form=..., time_order=k) integrators")
        self.integrators.append(integ)
        return self

    def Assemble(self, **_):
        fes = self.space
        while len(self._ops) < len(self.integrators):
            i = self.integrators[len(self._ops)]
            os_ = i.force_intorder if i.force_intorder >= 0 else None
            op = SpaceTimeOperator(fes, i.lset, i.dt, *i.coef, order_space=os_, order_time=i.time_order, share=self._ops[0] if self._ops else None)
            self._ops.append(op)
        for k, op in enumerate(self._ops):
            i = self.integrators[k]
            op.levelset, op.dt, op.coef = i.lset, i.dt, tuple(i.coef)
            op.Assemble(accumulate=k > 0)
        self.mat = self._ops[0].mat if self._ops else None
        return self
