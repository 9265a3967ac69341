#!/usr/bin/env python3
"""BASELINE configs[0] -- py_tutorials/fictdom_ghostpen.py at ITS order: 2D, [-1,1]^2, ring level set r1 = 1/4, r2 = 3/4, H1 order 3 with
dgjumps, operator = NEG stiffness + Nitsche boundary terms on IF + 0.1/h^2 ghost-penalty facet patches (fictdom_ghostpen.py:86-107), on the
undeformed mesh (order 3 on a deformed mesh is outside the CUDA path).  Structured triangles nx x nx x 2 (default nx = 700 -> 0.98 M).
One JSON line: Assemble time (CUDA events, 3 warm-ups), counts, clock record; beside it the CPU oracle (kind "port") on a bounded sample of
the patches.  A measurement for profiles/.   usage: tools/bench_fictdom2d.py [nx=700] [steps=10] [order=3]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    nx = int(sys.argv[1]) if len(sys.argv) > 1 else 700
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    order = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    import __graft_entry__ as ge
    ge.build()
    import bench
    from ngsxfem_b200 import meshes, xfem as xf, ngs
    dev = "cuda:0"
    mesh = ngs.Mesh(meshes.MakeStructured2DMesh(nx, lo=(-1, -1), hi=(1, 1), device=dev), device=dev)
    gf = ngs.GridFunction(ngs.H1(mesh, order=1))
    xf.InterpolateToP1(ngs.RingLevelset((0, 0), 0.25, 0.75), gf)
    ci = xf.CutInfo(mesh, gf)
    hasneg, hasif = ci.GetElementsOfType(xf.HASNEG), ci.GetElementsOfType(xf.IF)
    ba = xf.GetFacetsWithNeighborTypes(mesh, a=hasneg, b=hasif)
    Vh = ngs.H1(mesh, order=order, dgjumps=True)
    u, v = Vh.TnT()
    h, n = ngs.specialcf.mesh_size, ngs.LevelsetNormal(gf)
    lam = 10.0 * order * order
    a = xf.RestrictedBilinearForm(Vh, element_restriction=hasneg, facet_restriction=ba)
    a += xf.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xf.NEG}, form=ngs.grad(u) * ngs.grad(v), definedonelements=hasneg)
    a += xf.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xf.IF},              # Nitsche terms on the interface
                        form=-(ngs.grad(u) * n) * v - (ngs.grad(v) * n) * u + (lam / h) * u * v, definedonelements=hasif)
    a += xf.SymbolicFacetPatchBFI(form=0.1 * 1.0 / h * 1.0 / h * (u - u.Other()) * (v - v.Other()), skeleton=False, definedonelements=ba)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    clocks = bench.ClockSampler(0)
    clocks.start()
    t0 = time.time()
    a.Assemble()
    torch.cuda.synchronize()
    setup_s = time.time() - t0
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
    plan = a._plans[len(a.integrators) - 1]
    npatch = int(plan.els.shape[0])
    out = {"workload": f"configs[0] (fictdom_ghostpen.py) at H1 order {order}: 2D [-1,1]^2, ring level set, {mesh.ne} triangles, NEG stiffness"
                       + " + Nitsche IF terms + 0.1/h^2 ghost-penalty facet patches, undeformed mesh",
           "ne": mesh.ne, "hasneg_elements": int(hasneg.NumSet()), "if_elements": int(hasif.NumSet()), "ghost_facets": npatch, "ndof": int(Vh.ndof),
           "nnz": int(a.mat.nnz), "first_assemble_s(symbolic+numeric)": round(setup_s, 2), "assemble_ms": ms,
           "hasneg_elements_per_s": int(hasneg.NumSet()) / (ms * 1e-3), "patches_per_s": npatch / (ms * 1e-3)}
    try:
        from oracle import orc
        el2v, vert = mesh.el2v.cpu().numpy(), mesh.vert.cpu().numpy()
        pe = plan.els.cpu().numpy()
        sample = pe[::max(1, len(pe) // 4000)]
        t0 = time.perf_counter()
        for e0, e1 in sample:
            orc.facetpatch_elmat_h1(2, order, el2v[e0], vert[el2v[e0]], el2v[e1], vert[el2v[e1]], 0.1, -2)
        dt_s = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": len(sample) / dt_s, "unit": "patches/s", "cores": 1, "kind": "port",
                               "sample": f"{len(sample)} of the {len(pe)} patches, patch matrices only (orc_facetpatch_elmat_h1 incl. ctypes call overhead)"}
    except Exception as exc:
        out["cpu_baseline"] = {"error": f"{type(exc).__name__}: {str(exc)[:160]}"}
    out["clocks"] = clocks.stop()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
