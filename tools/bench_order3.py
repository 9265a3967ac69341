#!/usr/bin/env python3
"""Stage timing of H1 order 3 (k_elmat<3,3> dense blocks + k_rows_mats<20>): 3D Kuhn tets (default nx = 80 -> 3.07 M), sphere level set,
P3 H1 NEG stiffness + mass restricted to the HASNEG elements.  One JSON line; CUDA events, 3 warm-ups; beside it the CPU oracle
(kind "port": orc_elmat_mesh, OpenMP) on a bounded sample.  Not the bench.py contract -- a measurement for profiles/."""
import ctypes
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    nx = int(sys.argv[1]) if len(sys.argv) > 1 else 80
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    import __graft_entry__ as ge
    ge.build()
    from ngsxfem_b200 import _lib, meshes, ngs, xfem as xf
    L = _lib.load()
    mesh = ngs.Mesh(meshes.MakeStructured3DMesh(nx, device="cuda:0"), device="cuda:0")
    gf = ngs.GridFunction(ngs.H1(mesh, order=1))
    xf.InterpolateToP1(ngs.SphereLevelset((0.5, 0.5, 0.5), 0.35), gf)
    ci = xf.CutInfo(mesh, gf)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    import bench
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

    out = {"workload": f"H1 order 3: {mesh.ne} Kuhn tets, sphere r = 0.35, NEG grad-grad + mass restricted to HASNEG (order-6 rules: 64 points per "
                       "tet / sub-tet)", "ne": mesh.ne, "ms": {}}
    for fo in (2, 3):
        Vh = ngs.H1(mesh, order=fo)
        u, v = Vh.TnT()
        a = xf.RestrictedBilinearForm(Vh, element_restriction=ci.GetElementsOfType(xf.HASNEG))
        a += xf.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xf.NEG}, form=ngs.grad(u) * ngs.grad(v) + u * v)
        a.Assemble()
        assert a.assembly == "rows"
        plan = a._plans[0]
        n = plan.els.numel()
        out["hasneg_elements"] = n
        out[f"nnz_P{fo}"] = a.mat.nnz
        out["ms"][f"Assemble(P{fo})"] = timed(lambda: a.Assemble())
        if fo == 3:
            mc = ctypes.byref(mesh.c_struct())
            st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            mats = plan.mats
            dense = lambda: L.ngsx_elmat_h1(mc, gf.vec.data_ptr(), 3, 0, 6, ctypes.byref(a.integrators[0].form), n, plan.els.data_ptr(), mats.data_ptr(), None, st)
            out["ms"]["k_elmat3_fast + k_elmat3 (dense blocks)"] = timed(dense)
            L.ngsx_debug_set_order3_fma(1)
            out["ms"]["k_elmat<3,3> (first-generation FMA kernel, dense blocks)"] = timed(dense)
            L.ngsx_debug_set_order3_fma(0)
            dense()
            out["ms"]["k_rows_mats<20> (dense blocks)"] = timed(lambda: L.ngsx_assemble_csr_rows_mats(plan.handle, mats.data_ptr(), 0, a.mat.vals.data_ptr(), st))
            out["elements_per_s_P3_assemble"] = n / (out["ms"]["Assemble(P3)"] * 1e-3)
            out["producer_algorithmic_GBs"] = n * (16 + 96 + 32 + 3200) / (out["ms"]["k_elmat3_fast + k_elmat3 (dense blocks)"] * 1e-3) / 1e9
            try:
                from oracle import orc
                els = plan.els.cpu().numpy()
                sample = els[::max(1, len(els) // 100000)]
                el2v, vert, ls = mesh.el2v.cpu().numpy(), mesh.vert.cpu().numpy(), gf.vec.cpu().numpy()
                t0 = time.perf_counter()
                ref, _ = orc.elmat_mesh(3, el2v, vert, ls, 3, 0, 6, 1.0, 1.0, els=sample)
                dt_s = time.perf_counter() - t0
                got = plan.mats.reshape(-1, 20, 20)[::max(1, len(els) // 100000)].cpu().numpy()
                out["sample_parity_rel"] = float((np.abs(got - ref).reshape(len(ref), -1).max(1) / np.abs(ref).reshape(len(ref), -1).max(1)).max())
                out["cpu_baseline"] = {"value": len(sample) / dt_s, "unit": "elements/s", "cores": orc.load().orc_max_threads(), "kind": "port",
                                       "sample": f"{len(sample)} of the {len(els)} listed elements, element matrices only (orc_elmat_mesh, OpenMP)"}
            except Exception as exc:
                out["cpu_baseline"] = {"error": f"{type(exc).__name__}: {str(exc)[:200]}"}
    out["clocks"] = clocks.stop()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
