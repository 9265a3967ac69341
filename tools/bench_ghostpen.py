#!/usr/bin/env python3
"""Stage timings of the fictitious-domain operator at config-2 size:
3D Kuhn mesh nx (default 119), sphere level set, P2 H1 with dgjumps, NEG stiffness on HASNEG elements +
ghost-penalty facet patches on GetFacetsWithNeighborTypes(HASNEG, IF) (py_tutorials/fictdom_ghostpen.py:76-107).
Prints one JSON line; CUDA events, 3 warm-ups.  Not the bench.py contract -- a measurement for profiles/."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    nx = int(sys.argv[1]) if len(sys.argv) > 1 else 119
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    import __graft_entry__ as ge
    ge.build()
    from ngsxfem_b200 import meshes, xfem
    from ngsxfem_b200.ngs import GridFunction, H1, Mesh, SphereLevelset, grad, specialcf
    dev = "cuda:0"
    mesh = Mesh(meshes.MakeStructured3DMesh(nx, device=dev), device=dev)
    gf = GridFunction(H1(mesh, order=1))
    xfem.InterpolateToP1(SphereLevelset((0.5, 0.5, 0.5), 0.35), gf)
    ci = xfem.CutInfo(mesh, gf)
    hasneg, hasif = ci.GetElementsOfType(xfem.HASNEG), ci.GetElementsOfType(xfem.IF)
    mesh.facets()
    torch.cuda.synchronize()
    ev = lambda: torch.cuda.Event(enable_timing=True)
    import bench
    clocks = bench.ClockSampler(0)
    clocks.start()

    def timed(fn, n=steps, warm=3):
        for _ in range(warm):
            r = fn()
        torch.cuda.synchronize()
        a, b = ev(), ev()
        a.record()
        for _ in range(n):
            r = fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / n, r

    t_mark, ba = timed(lambda: xfem.GetFacetsWithNeighborTypes(mesh, a=hasneg, b=hasif))
    Vh = H1(mesh, order=2, dgjumps=True)
    u, v = Vh.TnT()
    h = specialcf.mesh_size
    a = xfem.RestrictedBilinearForm(Vh, element_restriction=hasneg, facet_restriction=ba)
    a += xfem.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xfem.NEG}, form=grad(u) * grad(v), definedonelements=hasneg)
    a += xfem.SymbolicFacetPatchBFI(form=0.1 * 1.0 / h * 1.0 / h * (u - u.Other()) * (v - v.Other()), skeleton=False, definedonelements=ba)
    t0 = time.time()
    a.Assemble()
    torch.cuda.synchronize()
    setup_s = time.time() - t0
    t_asm, _ = timed(lambda: a.Assemble())
    plan = a._plans[1]
    L = xfem._lib_()
    import ctypes
    from ngsxfem_b200.ngs import _ptr, _stream
    mc = ctypes.byref(mesh.c_struct())
    integ = a.integrators[1]
    def cmat():
        L.ngsx_set_packed_output(plan.slots)                       # packed block rows, as BilinearForm._assemble_facetpatch
        return L.ngsx_facetpatch_cmat_h1(mc, 2, integ.coef, integ.hpow, plan.els.shape[0], _ptr(plan.els), _ptr(plan.cidx), plan.ndp,
                                         _ptr(plan.rows), _stream())
    t_cmat, _ = timed(cmat)
    t_gather, _ = timed(lambda: L.ngsx_assemble_csr_rows_packed(plan.handle, _ptr(plan.rows), 1, _ptr(a.mat.vals), _stream()))
    npatch = plan.els.shape[0]
    out = {"workload": f"fictitious-domain operator, 3D Kuhn nx={nx} ({mesh.ne} tets), sphere r=0.35, P2 H1 dgjumps: NEG stiffness + ghost-penalty facet patches",
           "hasneg_elements": int(hasneg.NumSet()), "ghost_facets": npatch, "nnz": int(a.mat.nnz), "setup_s": round(setup_s, 2),
           "ms": {"GetFacetsWithNeighborTypes": t_mark, "Assemble (SetZero + volume + patches)": t_asm, "k_facetpatch_c (packed rows)": t_cmat, "k_rows_packed<14>": t_gather},
           "patches_per_s": npatch / ((t_cmat + t_gather) * 1e-3),
           "patch_algorithmic_GBs": npatch * (2 * (16 + 96) + 14 * 14 * 16) / ((t_cmat + t_gather) * 1e-3) / 1e9,
           "note": "algorithmic bytes per patch: two elements' connectivity + vertices (224 B) + 196 merged entries x (8 B read + 8 B write)"}
    out["clocks"] = clocks.stop()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
