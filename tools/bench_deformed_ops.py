#!/usr/bin/env python3
"""Stage timings of the operators of BASELINE configs 1 / 3 as the tutorials write them -- WITH mesh.SetDeformation -- at order 2, and
of the right-hand side f v on NEG at config-2 size (LinearForm through the per-dof gather):
  * fictitious-domain operator (NEG stiffness + 0.1/h^2 ghost-penalty facet patches, py_tutorials/fictdom_ghostpen.py:86-107) on the
    sphere geometry of config 2 with a synthetic order-2 deformation (what lsetcurving would hand over);
  * two-phase Nitsche-XFEM operator on FESpace([Vh, XFESpace(Vh)]) (py_tutorials/nxfem.py:77-107 / nxfem_higher_order.py) likewise;
  * LinearForm  (1 + x y) v  on NEG, affine and deformed.
One JSON line; CUDA events, 3 warm-ups, clock record.  Not the bench.py contract -- a measurement for profiles/.
usage: tools/bench_deformed_ops.py [nx=94] [steps=10]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def synthetic_deformation(ngs, mesh, nx):
    Vd = ngs.H1(mesh, order=2, dim=3)
    defo = ngs.GridFunction(Vd)
    h = 1.0 / nx
    X = mesh.vert
    d = 0.2 * h * torch.stack([torch.sin(7 * X[:, 1]) * torch.cos(5 * X[:, 2]), torch.sin(6 * X[:, 2] + X[:, 0]), torch.cos(8 * X[:, 0] * X[:, 1])], 1)
    defo.vec[:mesh.nv] = d
    g = torch.Generator(device="cpu").manual_seed(20241104)
    defo.vec[mesh.nv:] = (0.4 * h * (2 * torch.rand((Vd.ndof - mesh.nv, 3), generator=g, dtype=torch.float64) - 1)).to(mesh.device)
    return defo


def main():
    nx = int(sys.argv[1]) if len(sys.argv) > 1 else 94
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    import __graft_entry__ as ge
    ge.build()
    import bench
    from ngsxfem_b200 import meshes, xfem as xf, ngs
    from ngsxfem_b200.xfes import FESpace
    dev = "cuda:0"
    mesh = ngs.Mesh(meshes.MakeStructured3DMesh(nx, device=dev), device=dev)
    gf = ngs.GridFunction(ngs.H1(mesh, order=1))
    xf.InterpolateToP1(ngs.SphereLevelset((0.5, 0.5, 0.5), 0.35), gf)
    ci = xf.CutInfo(mesh, gf)
    hasneg, hasif = ci.GetElementsOfType(xf.HASNEG), ci.GetElementsOfType(xf.IF)
    defo = synthetic_deformation(ngs, mesh, nx)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    clocks = bench.ClockSampler(0)
    clocks.start()

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = ev(), ev()
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / steps

    out = {"workload": f"operators of configs 1 / 3 with mesh.SetDeformation (synthetic order-2 deformation), 3D Kuhn nx={nx} ({mesh.ne} tets), "
                       "sphere r = 0.35, P2 H1; right-hand side (1 + x y) v on NEG", "ne": mesh.ne, "hasneg_elements": int(hasneg.NumSet()),
           "if_elements": int(hasif.NumSet()), "ms": {}}
    h = ngs.specialcf.mesh_size

    def fictdom():
        ba = xf.GetFacetsWithNeighborTypes(mesh, a=hasneg, b=hasif)
        Vh = ngs.H1(mesh, order=2, dgjumps=True)
        u, v = Vh.TnT()
        a = xf.RestrictedBilinearForm(Vh, element_restriction=hasneg, facet_restriction=ba)
        a += xf.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xf.NEG}, form=ngs.grad(u) * ngs.grad(v), definedonelements=hasneg)
        a += xf.SymbolicFacetPatchBFI(form=0.1 * 1.0 / h * 1.0 / h * (u - u.Other()) * (v - v.Other()), skeleton=False, definedonelements=ba)
        return a

    def nxfem():
        Vh = ngs.H1(mesh, order=2)
        VhG = FESpace([Vh, xf.XFESpace(Vh, ci)])
        alpha = [1.0, 2.0]
        n = ngs.LevelsetNormal(gf)
        kappa = (xf.CutRatioGF(ci), 1.0 - xf.CutRatioGF(ci))
        stab = 20 * (alpha[1] + alpha[0]) / h
        (u_std, u_x), (v_std, v_x) = VhG.TrialFunction(), VhG.TestFunction()
        u = [u_std + op(u_x) for op in (xf.neg, xf.pos)]
        v = [v_std + op(v_x) for op in (xf.neg, xf.pos)]
        gradu = [ngs.grad(u_std) + op(u_x) for op in (xf.neg_grad, xf.pos_grad)]
        gradv = [ngs.grad(v_std) + op(v_x) for op in (xf.neg_grad, xf.pos_grad)]
        fu = sum([-kappa[i] * alpha[i] * gradu[i] * n for i in (0, 1)])
        fv = sum([-kappa[i] * alpha[i] * gradv[i] * n for i in (0, 1)])
        a = xf.BilinearForm(VhG)
        a += xf.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xf.NEG}, form=alpha[0] * gradu[0] * gradv[0])
        a += xf.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xf.POS}, form=alpha[1] * gradu[1] * gradv[1])
        a += xf.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xf.IF},
                            form=fu * (v[0] - v[1]) + fv * (u[0] - u[1]) + stab * (u[0] - u[1]) * (v[0] - v[1]))
        return a

    def rhs():
        Vh = ngs.H1(mesh, order=2)
        v = Vh.TestFunction()
        f = xf.LinearForm(Vh)
        f += xf.SymbolicLFI(levelset_domain={"levelset": gf, "domain_type": xf.NEG}, form=(1.0 + ngs.x * ngs.y) * v)
        return f

    for name, make in (("fictdom operator (NEG stiffness + ghost penalty)", fictdom), ("Nitsche-XFEM operator (NEG + POS + IF)", nxfem),
                       ("LinearForm (1 + x y) v on NEG", rhs)):
        for geom in ("affine", "deformed"):
            if geom == "deformed":
                mesh.SetDeformation(defo)
            else:
                mesh.UnsetDeformation()
            try:
                obj = make()
                t0 = time.time()
                obj.Assemble()
                torch.cuda.synchronize()
                out["ms"][f"{name}, {geom}"] = timed(lambda: obj.Assemble())
                out.setdefault("setup_s", {})[f"{name}, {geom}"] = round(time.time() - t0, 2)
                if hasattr(obj, "mat"):
                    out.setdefault("nnz", {})[name] = int(obj.mat.nnz)
                del obj
            except Exception as exc:                                   # a path outside the CUDA path: recorded, not hidden
                out["ms"][f"{name}, {geom}"] = f"{type(exc).__name__}: {str(exc)[:160]}"
            torch.cuda.empty_cache()
    mesh.UnsetDeformation()
    nh = out["hasneg_elements"]
    t = out["ms"].get("LinearForm (1 + x y) v on NEG, affine")
    if isinstance(t, float):
        out["rhs_elements_per_s_affine"] = nh / (t * 1e-3)
        out["rhs_algorithmic_GBs_affine"] = nh * (16 + 96 + 32 + 40 + 80) / (t * 1e-3) / 1e9      # el2v + vertices + level set + dof numbers + 10 entries written
    out["clocks"] = clocks.stop()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
