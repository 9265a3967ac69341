#!/usr/bin/env python3
"""Stage timing of BASELINE.json config 4: TraceFEM Laplace-Beltrami operator on a sphere (py_tutorials/tracefem3d.py:70-88
forms, undeformed, order 2) + ghost-penalty facet patches between IF elements, 3D Kuhn mesh nx (default 94 -> 4.98 M tets).
One JSON line; CUDA events, 3 warm-ups.  Not the bench.py contract -- a measurement for profiles/."""
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
    dev = "cuda:0"
    mesh = ngs.Mesh(meshes.MakeStructured3DMesh(nx, device=dev), device=dev)
    gf = ngs.GridFunction(ngs.H1(mesh, order=1))
    xf.InterpolateToP1(ngs.SphereLevelset((0.5, 0.5, 0.5), 0.35), gf)
    ci = xf.CutInfo(mesh, gf)
    ba_if = ci.GetElementsOfType(xf.IF)
    ba_f = xf.GetFacetsWithNeighborTypes(mesh, a=ba_if, b=ba_if)
    Vh = ngs.H1(mesh, order=2, dgjumps=True)
    u, v = Vh.TnT()
    n, h = ngs.LevelsetNormal(gf), ngs.specialcf.mesh_size
    P = lambda w: w - (w * n) * n
    a = xf.RestrictedBilinearForm(Vh, element_restriction=ba_if, facet_restriction=ba_f)
    a += xf.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xf.IF}, form=P(ngs.grad(u)) * P(ngs.grad(v)) + u * v, definedonelements=ba_if)
    a += xf.SymbolicFacetPatchBFI(form=0.1 / h / h * (u - u.Other()) * (v - v.Other()), skeleton=False, definedonelements=ba_f)
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
    nif, npatch = int(ba_if.NumSet()), int(a._plans[1].els.shape[0])
    out = {"workload": f"configs[3]: TraceFEM Laplace-Beltrami operator (IF mass + tangential stiffness) + ghost-penalty facet patches, "
                       f"P2, 3D Kuhn nx={nx} ({mesh.ne} tets), sphere r=0.35, undeformed",
           "ne": mesh.ne, "if_elements": nif, "ghost_facets": npatch, "nnz": int(a.mat.nnz),
           "first_assemble_s(symbolic+numeric)": round(setup_s, 2), "assemble_ms": ms,
           "if_elements_per_s": nif / (ms * 1e-3), "patches_per_s": npatch / (ms * 1e-3)}
    out["clocks"] = clocks.stop()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
