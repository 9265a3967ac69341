#!/usr/bin/env python3
"""Stage timing of BASELINE.json config 5: space-time cut quadrature of a moving circle, 2D + t, order 2 in space x order 2 in
time, nx x nx x 2 triangles (default nx = 1000 -> 2 M).  One JSON line; CUDA events, 3 warm-ups; next to it the compiled
reference (oracle/_ref: spacetimecutrule.cpp) timed single-threaded on a bounded sample of the same elements.
Not the bench.py contract -- a measurement for profiles/."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    nx = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    import __graft_entry__ as ge
    ge.build()
    from ngsxfem_b200 import meshes, xfem as xf, ngs, spacetime
    mesh = ngs.Mesh(meshes.MakeStructured2DMesh(nx, device="cuda:0"), device="cuda:0")
    gf = ngs.GridFunction(xf.SpaceTimeFESpace(ngs.H1(mesh, order=1), xf.ScalarTimeFE(2)))
    xf.SpaceTimeInterpolateToP1(lambda xx, yy, t: torch.sqrt((xx - 0.35 - 0.3 * t) ** 2 + (yy - 0.5) ** 2) - 0.25, None, gf)
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

    out = {"workload": f"configs[4]: space-time cut quadrature, moving circle, 2D+t, {mesh.ne} triangles, order 2 (space) x 2 (time), "
                       "level set P1 x nodal order 2 in time", "ne": mesh.ne, "ms": {}}
    for name, dt in (("NEG", xf.NEG), ("IF", xf.IF)):
        t_int, val = timed(lambda: xf.IntegrateX(gf, mesh, cf=1.0, order=2, domain_type=dt, time_order=2, sync=False))
        t_rule, rule = timed(lambda: spacetime.cut_rule(gf, mesh, dt, 2, 2))
        out["ms"][f"Integrate({name})"] = t_int
        out["ms"][f"rule count+scan+fill({name})"] = t_rule
        out[f"points_{name}"] = int(rule[2].shape[0])
        out[f"value_{name}"] = float(val.item())
    # f-3 remainder: space-time CutInfo, SpaceTimeFE element matrices (P1 x nodal order 1: 6 x 6 blocks) and the assembled operator
    out["ms"]["CutInfo.Update(time_order=2)"], _ = timed(lambda: spacetime.classify(gf, mesh, 2))
    fes = xf.SpaceTimeFESpace(ngs.H1(mesh, order=1), xf.ScalarTimeFE(1))
    blocks = torch.empty((mesh.ne, fes.nd, fes.nd), dtype=torch.float64, device="cuda:0")
    out["ms"]["SpaceTimeFE element matrices (NEG, mass + stiffness + dt)"], _ = timed(lambda: spacetime.elmats(fes, gf, xf.NEG, 1.0, 1.0, 1.0, out=blocks))
    del blocks
    op = spacetime.SpaceTimeOperator(fes, gf, xf.NEG, 1.0, 1.0, 1.0)
    out["ms"]["SpaceTimeOperator.Assemble (blocks + row gather)"], _ = timed(lambda: op.Assemble())
    out["nnz_spacetime_operator"] = op.mat.nnz
    out["elements_per_s_st_assemble"] = mesh.ne / (out["ms"]["SpaceTimeOperator.Assemble (blocks + row gather)"] * 1e-3)
    out["elements_per_s_integrate_NEG"] = mesh.ne / (out["ms"]["Integrate(NEG)"] * 1e-3)
    out["points_per_s_fill_NEG"] = out["points_NEG"] / (out["ms"]["rule count+scan+fill(NEG)"] * 1e-3)
    try:
        from oracle import orc, refload
        if refload.available():
            el2v, vert = mesh.el2v.cpu().numpy(), mesh.vert.cpu().numpy()
            ls = gf.vec.cpu().numpy().reshape(3, mesh.nv)
            sample = np.arange(0, mesh.ne, max(1, mesh.ne // 20000))
            t0 = time.perf_counter()
            for e in sample:
                orc.st_cut_rule(2, vert[el2v[e]], ls[:, el2v[e]], 0, 2, 2, impl="ref")
            dt_s = time.perf_counter() - t0
            out["cpu_reference"] = {"elements_per_s": len(sample) / dt_s, "cores": 1, "kind": "reference",
                                    "sample": f"{len(sample)} elements through oracle/_ref (spacetimecutrule.cpp) incl. ctypes call overhead"}
    except Exception as exc:
        out["cpu_reference"] = {"error": str(exc)[:100]}
    out["clocks"] = clocks.stop()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
