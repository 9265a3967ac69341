#!/usr/bin/env python3
"""BASELINE config 2 with symmetric storage (BilinearForm(symmetric=True): SparseMatrixSymmetric keeps the lower triangle,
NGS/linalg/sparsematrix_impl.hpp:605-690) beside the unsymmetric headline: Assemble time, stored values, and the device -> host
download that bounds the end-to-end step.  One JSON line; CUDA events, 3 warm-ups, clock record.  A measurement for profiles/.
usage: tools/bench_symmetric.py [nx=119] [steps=10]"""
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
    import bench
    from ngsxfem_b200 import meshes, xfem as xf, ngs
    dev = "cuda:0"
    mesh = ngs.Mesh(meshes.MakeStructured3DMesh(nx, device=dev), device=dev)
    gf = ngs.GridFunction(ngs.H1(mesh, order=1))
    xf.InterpolateToP1(ngs.SphereLevelset((0.5, 0.5, 0.5), 0.35), gf)
    ci = xf.CutInfo(mesh, gf)
    hasneg = ci.GetElementsOfType(xf.HASNEG)
    Vh = ngs.H1(mesh, order=2)
    u, v = Vh.TnT()
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

    out = {"workload": f"configs[1] with symmetric storage beside the unsymmetric headline: 3D Kuhn nx={nx} ({mesh.ne} tets), sphere r = 0.35, P2 H1, "
                       "NEG grad-grad + mass restricted to HASNEG", "hasneg_elements": int(hasneg.NumSet()), "ms": {}, "nnz": {}, "d2h_ms": {}}
    vals = {}
    for name, sym in (("unsymmetric", False), ("symmetric", True)):
        a = xf.RestrictedBilinearForm(Vh, element_restriction=hasneg, symmetric=sym)
        a += xf.SymbolicBFI(levelset_domain={"levelset": gf, "domain_type": xf.NEG}, form=ngs.grad(u) * ngs.grad(v) + u * v)
        t0 = time.time()
        a.Assemble()
        torch.cuda.synchronize()
        out.setdefault("setup_s", {})[name] = round(time.time() - t0, 2)
        out["ms"][f"Assemble ({name})"] = timed(lambda: a.Assemble())
        out["nnz"][name] = int(a.mat.nnz)
        host = torch.empty(a.mat.nnz, dtype=torch.float64).pin_memory()
        out["d2h_ms"][name] = timed(lambda: host.copy_(a.mat.vals, non_blocking=True))
        vals[name] = (a.mat.rowptr.clone(), a.mat.colidx.clone(), a.mat.vals.clone())
        del a, host
    rp, cidx, va = vals["unsymmetric"]
    rows = torch.repeat_interleave(torch.arange(Vh.ndof, device=dev), rp[1:] - rp[:-1])
    keep = cidx.long() <= rows
    out["lower_triangle_bitwise_equal"] = bool(torch.equal(va[keep], vals["symmetric"][2]) and torch.equal(cidx[keep], vals["symmetric"][1]))
    out["e2e_bound_ms"] = {k: out["ms"][f"Assemble ({k})"] + 0.0 if False else max(out["d2h_ms"][k], out["ms"][f"Assemble ({k})"]) for k in ("unsymmetric", "symmetric")}
    out["note"] = "e2e_bound_ms = max(download, Assemble): with the double-buffered pipeline of bench.py the step is bounded by the slower of the two"
    out["clocks"] = clocks.stop()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
