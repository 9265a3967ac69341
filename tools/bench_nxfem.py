#!/usr/bin/env python3
"""Stage timings of BASELINE.json config 3: two-phase Nitsche-XFEM operator (py_tutorials/nxfem.py:62-107 forms), order 2,
3D Kuhn mesh nx (default 94 -> 4.98 M tets), sphere level set.  One JSON line; CUDA events, 3 warm-ups.
Not the bench.py contract -- a measurement for profiles/."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    nx = int(sys.argv[1]) if len(sys.argv) > 1 else 94
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    import __graft_entry__ as ge
    ge.build()
    from ngsxfem_b200 import meshes, xfem as xf, ngs
    from ngsxfem_b200.xfes import FESpace
    dev = "cuda:0"
    mesh = ngs.Mesh(meshes.MakeStructured3DMesh(nx, device=dev), device=dev)
    gf = ngs.GridFunction(ngs.H1(mesh, order=1))
    xf.InterpolateToP1(ngs.SphereLevelset((0.5, 0.5, 0.5), 0.35), gf)
    ci = xf.CutInfo(mesh, gf)
    Vh = ngs.H1(mesh, order=2)
    t0 = time.time()
    Vhx = xf.XFESpace(Vh, ci)
    torch.cuda.synchronize()
    t_xfes = time.time() - t0
    VhG = FESpace([Vh, Vhx])
    alpha = [1.0, 2.0]
    n, h = ngs.LevelsetNormal(gf), ngs.specialcf.mesh_size
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
    t0 = time.time()
    a.Assemble()
    torch.cuda.synchronize()
    setup_s = time.time() - t0
    ev = lambda: torch.cuda.Event(enable_timing=True)
    import bench
    clocks = bench.ClockSampler(0)
    clocks.start()
    for _ in range(3):
        a.Assemble()
    torch.cuda.synchronize()
    b, e = ev(), ev()
    b.record()
    for _ in range(steps):
        a.Assemble()
    e.record()
    torch.cuda.synchronize()
    ms = b.elapsed_time(e) / steps
    a.check_degenerate()
    sym = (a.mat.to_scipy() - a.mat.to_scipy().T)
    out = {"workload": f"configs[2]: two-phase Nitsche-XFEM operator, order 2, 3D Kuhn nx={nx} ({mesh.ne} tets), sphere r=0.35, "
                       "FESpace([H1, XFESpace]): NEG + POS stiffness with neg_grad/pos_grad, IF Nitsche terms with CutRatioGF averages",
           "ne": mesh.ne, "if_elements": int(Vhx.if_els.numel()), "ndof": VhG.ndof, "x_ndof": Vhx.ndof, "nnz": int(a.mat.nnz),
           "xfespace_update_s": round(t_xfes, 3), "first_assemble_s(symbolic+numeric)": round(setup_s, 2),
           "assemble_ms": ms, "elements_per_s": mesh.ne / (ms * 1e-3),
           "asymmetry_rel": float(abs(sym).max() / abs(a.mat.to_scipy()).max())}
    out["clocks"] = clocks.stop()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
