#!/usr/bin/env python3
"""Stage timing of SURVEY section 8 row f-4 at BASELINE config-2 size: 3D Kuhn tets (default nx = 119 -> 10.1 M), sphere level set,
P2 H1 NEG stiffness + mass on a DEFORMED mesh (order-2 vector H1 deformation, smooth field of amplitude 0.2 h), restricted to
the HASNEG elements, plus Integrate(IF) / Integrate(NEG) on the deformed mesh.  One JSON line; CUDA events, 3 warm-ups; next to
it the CPU oracle (kind "port": oracle/ngsx_oracle.c orc_elmat_mesh_iso, OpenMP) on a bounded sample of the listed elements.
Not the bench.py contract -- a measurement for profiles/."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    nx = int(sys.argv[1]) if len(sys.argv) > 1 else 119
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    dord = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    import __graft_entry__ as ge
    ge.build()
    from ngsxfem_b200 import _lib, meshes, ngs, xfem as xf
    L = _lib.load()
    mesh = ngs.Mesh(meshes.MakeStructured3DMesh(nx, device="cuda:0"), device="cuda:0")
    gf = ngs.GridFunction(ngs.H1(mesh, order=1))
    xf.InterpolateToP1(ngs.SphereLevelset((0.5, 0.5, 0.5), 0.35), gf)
    ci = xf.CutInfo(mesh, gf)
    Vd = ngs.H1(mesh, order=dord, dim=3)
    defo = ngs.GridFunction(Vd)
    h = 1.0 / nx
    X = mesh.vert
    d = 0.2 * h * torch.stack([torch.sin(7 * X[:, 1]) * torch.cos(5 * X[:, 2]), torch.sin(6 * X[:, 2] + X[:, 0]), torch.cos(8 * X[:, 0] * X[:, 1])], 1)
    defo.vec[:mesh.nv] = d
    if dord == 2:
        g = torch.Generator(device="cpu").manual_seed(20241104)
        defo.vec[mesh.nv:] = (0.4 * h * (2 * torch.rand((Vd.ndof - mesh.nv, 3), generator=g, dtype=torch.float64) - 1)).to(mesh.device)
    Vh = ngs.H1(mesh, order=2)
    u, v = Vh.TnT()
    form = ngs.grad(u) * ngs.grad(v) + u * v
    ev = lambda: torch.cuda.Event(enable_timing=True)
    import bench
    clocks = bench.ClockSampler(0)
    clocks.start()

    def timed(fn):
        for _ in range(3):
            r = fn()
        torch.cuda.synchronize()
        a, b = ev(), ev()
        a.record()
        for _ in range(steps):
            r = fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / steps, r

    out = {"workload": f"f-4: configs[1] geometry ({mesh.ne} tets, sphere r = 0.35) on a deformed mesh (vector H1 order {dord}), P2 H1 NEG "
                       "grad-grad + mass, restricted to HASNEG; Integrate(IF), Integrate(NEG) on the deformed mesh", "ne": mesh.ne, "ms": {}}
    res = {}
    for name, deformed in (("affine", False), ("deformed", True)):
        if deformed:
            mesh.SetDeformation(defo)
        else:
            mesh.UnsetDeformation()
        a = xf.RestrictedBilinearForm(Vh, element_restriction=ci.GetElementsOfType(xf.HASNEG))
        a += xf.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xf.NEG}, form=form)
        a.Assemble()
        n = a._plans[0].els.numel()
        out["hasneg_elements"] = n
        t_asm, _ = timed(lambda: a.Assemble())
        out["ms"][f"Assemble({name})"] = t_asm
        if deformed:
            plan = a._plans[0]
            lv = gf.vec
            dd = mesh.c_deformation()
            import ctypes
            mc = ctypes.byref(mesh.c_struct())
            st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            def elmat_packed():
                L.ngsx_set_packed_output(plan.slots)
                return L.ngsx_elmat_h1_deformed(mc, ctypes.byref(dd), lv.data_ptr(), 2, 0, 4, ctypes.byref(a.integrators[0].form), n,
                                                plan.els.data_ptr(), plan.rows.data_ptr(), None, None, st)
            t_el, _ = timed(elmat_packed)
            t_ga, _ = timed(lambda: L.ngsx_assemble_csr_rows_packed(plan.handle, plan.rows.data_ptr(), 0, a.mat.vals.data_ptr(), st))
            out["ms"]["k_elmat_def (packed rows)"] = t_el
            out["ms"]["k_rows_packed"] = t_ga
            dense = torch.empty((n, 100), dtype=torch.float64, device=mesh.device)
            t_eld, _ = timed(lambda: L.ngsx_elmat_h1_deformed(mc, ctypes.byref(dd), lv.data_ptr(), 2, 0, 4, ctypes.byref(a.integrators[0].form), n,
                                                              plan.els.data_ptr(), dense.data_ptr(), None, None, st))
            t_gad, _ = timed(lambda: L.ngsx_assemble_csr_rows_mats(plan.handle, dense.data_ptr(), 0, a.mat.vals.data_ptr(), st))
            out["ms"]["k_elmat_def (dense blocks)"] = t_eld
            out["ms"]["k_rows_mats (dense blocks)"] = t_gad
            out["pairs"] = int(plan.npairs)
            out["elements_per_s_deformed_assemble"] = n / (t_asm * 1e-3)
            # algorithmic bytes of the deformed element-matrix kernel: 16 B el2v + 40 B deformation dofs + 10 x 24 B coefficients + 96 B
            # vertices + 32 B level set in, 800 B block out
            out["k_elmat_def_algorithmic_GBs"] = n * (16 + 40 + 240 + 96 + 32 + 800) / (t_el * 1e-3) / 1e9
        for dname, dt in (("IF", xf.IF), ("NEG", xf.NEG)):
            t_i, val = timed(lambda: xf.IntegrateX(gf, mesh, cf=1.0, order=4, domain_type=dt, sync=False))
            out["ms"][f"Integrate({dname},{name})"] = t_i
            res[(dname, name)] = float(val.item())
        res[name] = a.mat.vals.clone()
    out["values"] = {f"{k[0]}_{k[1]}": v for k, v in res.items() if isinstance(k, tuple)}
    out["csr_rel_change_by_deformation"] = float(((res["deformed"] - res["affine"]).abs().max() / res["affine"].abs().max()).item())
    try:
        from oracle import orc
        els = a._plans[0].els.cpu().numpy()
        sample = els[::max(1, len(els) // 200000)]
        el2v, vert, ls = mesh.el2v.cpu().numpy(), mesh.vert.cpu().numpy(), gf.vec.cpu().numpy()
        e2dd, dv = Vd.el2dof.cpu().numpy(), defo.vec.cpu().numpy()
        t0 = time.perf_counter()
        ref, _ = orc.elmat_mesh_iso(3, el2v, vert, ls, dord, e2dd, dv, 2, 0, 4, orc.form8(1.0, 1.0), els=sample)
        dt_s = time.perf_counter() - t0
        got = dense.reshape(-1, 10, 10)[::max(1, len(els) // 200000)].cpu().numpy()
        out["sample_parity_rel"] = float((np.abs(got - ref).reshape(len(ref), -1).max(1) / np.abs(ref).reshape(len(ref), -1).max(1)).max())
        out["cpu_baseline"] = {"value": len(sample) / dt_s, "unit": "elements/s", "cores": orc.load().orc_max_threads(), "kind": "port",
                               "sample": f"{len(sample)} of the {len(els)} listed elements, element matrices only (orc_elmat_mesh_iso, OpenMP)"}
    except Exception as exc:
        out["cpu_baseline"] = {"error": f"{type(exc).__name__}: {str(exc)[:200]}"}
    out["clocks"] = clocks.stop()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
