"""Space-time level sets and cut quadrature on the CUDA path: stand-ins for ScalarTimeFE,
SpaceTimeFESpace and SpaceTimeInterpolateToP1 (spacetime/python_spacetime.cpp, spacetime/SpaceTimeFESpace.cpp) as far as the
space-time cut rule needs them -- a level set that is P1 in space and nodal in the reference time.  The quadrature itself is
ngsx_st_rule_count / _fill / ngsx_st_integrate (cutint/spacetimecutrule.cpp:98-196)."""
import ctypes

import torch

from . import _lib
from ._lib import NgsxError, check
from .ngs import FESpace, GridFunction, H1, Poly, TPoly, _ptr, _stream

_NODES = {0: [0.0], 1: [0.0, 1.0], 2: [0.0, 0.5, 1.0]}          # NodalTimeFE::CalcInterpolationPoints, spacetime/SpaceTimeFE.cpp:150-160


class ScalarTimeFE:
    """spacetime/python_spacetime.cpp: nodal time element of the given order (orders 0..2 on the CUDA path)"""

    def __init__(self, order: int = 1, skip_first_node: bool = False, only_first_node: bool = False):
        if skip_first_node or only_first_node or order not in _NODES:
            raise NgsxError(-2, "ScalarTimeFE: orders 0..2 without node skipping are on the CUDA path (order > 2 needs the bisection "
                                "root finding of cutint/spacetimecutrule.cpp:46-95)")
        self.order = order
        self.nodes = _NODES[order]
        self.ndof = order + 1


class SpaceTimeFESpace(FESpace):
    """tensor-product space: dof = k * ndof_space + space dof (spacetime/SpaceTimeFESpace.cpp:109-122)"""

    def __init__(self, space: H1, tfe: ScalarTimeFE, **_):
        if not isinstance(space, H1) or space.order != 1:
            raise NgsxError(-2, "SpaceTimeFESpace: the space-time level set must be P1 in space")
        self.space, self.tfe, self.mesh, self.order = space, tfe, space.mesh, space.order
        self.ndof = space.ndof * tfe.ndof

    def TimeFE_nodes(self):
        return list(self.tfe.nodes)


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
