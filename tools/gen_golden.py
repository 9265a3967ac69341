#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the COMPILED REFERENCE (oracle/_ref/libngsxref.so = the reference's own
straightcutrule.cpp + NGSolve fem sources, built in place by oracle/build_ref.sh).

Run in the dev container only (the reference tree does not travel to the GPU box):
    python tools/gen_golden.py
The fixtures are small, seeded and committed; tests compare the oracle (not-gpu) and the CUDA path (gpu) to them.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import orc, refload  # noqa: E402
from ngsxfem_b200 import meshes  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
SEED = 20241104


def nice_simplex(rng, dim):
    pm = np.vstack([np.eye(dim), np.zeros(dim)]) + 0.15 * rng.uniform(-1, 1, (dim + 1, dim))
    return pm * rng.uniform(0.05, 0.5) + rng.uniform(-1, 1, dim)


def gen_elements(dim, n, rng):
    """seeded single elements incl. the degenerate families of SURVEY.md section 3.1"""
    pm = np.stack([nice_simplex(rng, dim) for _ in range(n)])
    ls = rng.uniform(-1, 1, (n, dim + 1))
    for t in range(n):
        if t % 9 == 1: ls[t] = np.abs(ls[t])
        if t % 9 == 2: ls[t] = -np.abs(ls[t])
        if t % 9 == 3: ls[t, rng.integers(dim + 1)] = 1e-14          # InterpolateToP1 clamp value
        if t % 9 == 4: ls[t] = np.abs(ls[t]); ls[t, rng.integers(dim + 1)] = 0.0   # zero + all positive -> POS
        if t % 9 == 5: ls[t] = -np.abs(ls[t]); ls[t, rng.integers(dim + 1)] = 0.0  # zero + all negative -> NEG
        if t % 27 == 6: ls[t, 0] = 0.0; ls[t, 1] = 1.0; ls[t, 2] = -1.0            # strict mixed + exact zero -> reference throws
        if t % 27 == 7: ls[t] = 0.0                                                # all zero -> reference throws
        if t % 27 == 8: ls[t, 0] = -0.0; ls[t, 1:] = np.abs(ls[t, 1:])
    vn = np.stack([rng.permutation(1000)[:dim + 1] for _ in range(n)]).astype(np.int32)
    return pm, ls, vn


def single_elements():
    rng = np.random.default_rng(SEED)
    out = {}
    for dim in (2, 3):
        n = 120
        pm, ls, vn = gen_elements(dim, n, rng)
        out[f"d{dim}_pmat"] = pm; out[f"d{dim}_lset"] = ls; out[f"d{dim}_vnums"] = vn
        dom = np.zeros(n, dtype=np.int8); ratio = np.zeros(n)
        for t in range(n):
            dom[t], ratio[t] = orc.cutinfo_element(dim, pm[t], ls[t], impl="ref")
        out[f"d{dim}_dom"] = dom; out[f"d{dim}_ratio"] = ratio
        for dt in (0, 1, 2):
            for order in (0, 1, 2, 3, 4, 5, 6):
                npts = np.zeros(n, dtype=np.int32); P = []; W = []
                for t in range(n):
                    k, p, w = orc.cut_rule(dim, pm[t], ls[t], dt, order, impl="ref")
                    npts[t] = k
                    if k > 0:
                        P.append(p); W.append(w)
                out[f"d{dim}_dt{dt}_o{order}_npts"] = npts
                out[f"d{dim}_dt{dt}_o{order}_pts"] = np.concatenate(P) if P else np.zeros((0, 3))
                out[f"d{dim}_dt{dt}_o{order}_wts"] = np.concatenate(W) if W else np.zeros(0)
        for fo in (1, 2, 3):
            nd = orc.load().orc_h1_ndof(dim, fo)
            for dt in (0, 1, 2):
                em = np.zeros((n, nd, nd)); ev = np.zeros((n, nd)); npt = np.zeros(n, dtype=np.int32)
                for t in range(n):
                    npt[t], m = orc.elmat_h1(dim, fo, vn[t], pm[t], ls[t], dt, 2 * fo, 1.25, 0.75, impl="ref")
                    if npt[t] >= 0: em[t] = m
                    k, v = orc.elvec_h1(dim, fo, vn[t], pm[t], ls[t], dt, 2 * fo, [(1.0, 0, 0, 0), (2.0, 1, 2, 0), (-0.5, 0, 0, 1)], impl="ref")
                    if k >= 0: ev[t] = v
                out[f"d{dim}_p{fo}_dt{dt}_elmat"] = em; out[f"d{dim}_p{fo}_dt{dt}_elvec"] = ev
                out[f"d{dim}_p{fo}_dt{dt}_npts"] = npt
    np.savez_compressed(os.path.join(OUT, "single_elements.npz"), **out)
    print("single_elements.npz", len(out), "arrays")


def small_meshes():
    """config-2-shaped mesh (Kuhn cube, sphere r=0.35) at nx=6, and the 2D ring of fictdom_ghostpen at nx=12"""
    out = {}
    m3 = meshes.MakeStructured3DMesh(6)
    ls3 = meshes.sphere_levelset(m3.vert, (0.5, 0.5, 0.5), 0.35).numpy()
    ls3 = np.where(np.abs(ls3) < 1e-14, 1e-14, ls3)
    m2 = meshes.MakeStructured2DMesh(12, lo=(-1, -1), hi=(1, 1))
    r = np.sqrt((m2.vert.numpy() ** 2).sum(1))
    ls2 = np.maximum(0.25 - r, r - 0.75)   # ring |r-rc|-rr, r1=1/4, r2=3/4 (fictdom_ghostpen.py:28-33)
    ls2 = np.where(np.abs(ls2) < 1e-14, 1e-14, ls2)
    for tag, m, ls in (("m3", m3, ls3), ("m2", m2, ls2)):
        dim = m.dim
        el2v = m.el2v.numpy(); vert = m.vert.numpy()
        out[f"{tag}_lset"] = ls
        dom, ratio, nerr = orc.cutinfo_mesh(dim, el2v, vert, ls, impl="ref")
        assert nerr == 0
        out[f"{tag}_dom"] = dom; out[f"{tag}_ratio"] = ratio
        for dt in (0, 1, 2):
            for order in (0, 2, 4):
                v, npt = orc.integrate_mesh(dim, el2v, vert, ls, dt, order, [(1.0, 0, 0, 0)], impl="ref", nthreads=1)
                v2, _ = orc.integrate_mesh(dim, el2v, vert, ls, dt, order, [(1.0, 2, 0, 0), (0.5, 0, 1, 1 if dim == 3 else 0)], impl="ref", nthreads=1)
                out[f"{tag}_int_dt{dt}_o{order}"] = np.array([v, v2, npt])
            for fo in (1, 2):
                em, npts = orc.elmat_mesh(dim, el2v, vert, ls, fo, dt, 2 * fo, 1.0, 1.0, impl="ref")
                out[f"{tag}_p{fo}_dt{dt}_elmat"] = em; out[f"{tag}_p{fo}_dt{dt}_npts"] = npts
    np.savez_compressed(os.path.join(OUT, "small_meshes.npz"), **out)
    print("small_meshes.npz", len(out), "arrays")


def ref_rule_et(et, ls, dt, order, cap=8192):
    """StraightCutIntegrationRuleUntransformed of the compiled reference for any element type (ET_QUAD = 11, ET_HEX = 24)"""
    import ctypes
    dp = ctypes.POINTER(ctypes.c_double)
    ls = np.ascontiguousarray(ls, dtype=np.float64); pts = np.zeros((cap, 3)); w = np.zeros(cap)
    n = refload.load().ref_cut_rule_untrafo(et, ls.ctypes.data_as(dp), dt, order, pts.ctypes.data_as(dp), w.ctypes.data_as(dp), cap)
    return n, pts[:max(n, 0)].copy(), w[:max(n, 0)].copy()


def quad_hex_rules(only_return=False):
    """SURVEY section 8 row f-5 groundwork (no CUDA path yet): cut rules of the reference's LevelsetCutQuadrilateral
    (cutint/straightcutrule.cpp:206-501) on the reference square / cube for seeded multilinear level sets (vertex values in
    NGSolve's vertex order), FIND_OPTIMAL policy -- the target a quad / hex kernel has to meet.  Ragged rules are stored flat
    with offsets; n = -1: the reference returns no rule (element outside the domain), -2: it throws."""
    rng = np.random.default_rng(SEED + 5)
    out = {}
    for name, et, nv, n in (("quad", 11, 4, 80), ("hex", 24, 8, 60)):
        ls = rng.uniform(-1, 1, (n, nv))
        for t in range(n):
            if t % 8 == 1: ls[t] = np.abs(ls[t])
            if t % 8 == 2: ls[t] = -np.abs(ls[t])
            if t % 8 == 3: ls[t] = rng.uniform(-1, 1) + rng.uniform(-1, 1, 3)[:2 if nv == 4 else 3] @ (np.array(
                [[0, 1, 1, 0], [0, 0, 1, 1]] if nv == 4 else [[0, 1, 1, 0, 0, 1, 1, 0], [0, 0, 1, 1, 0, 0, 1, 1], [0, 0, 0, 0, 1, 1, 1, 1]], dtype=float))   # affine
            if t % 8 == 4: ls[t, rng.integers(nv)] = 1e-14
        out[f"{name}_lset"] = ls
        for dt in (0, 1, 2):
            for order in (1, 2, 4):
                cnt = np.zeros(n, dtype=np.int32); P = []; W = []
                for t in range(n):
                    k, p, w = ref_rule_et(et, ls[t], dt, order)
                    cnt[t] = k
                    if k > 0:
                        P.append(p); W.append(w)
                out[f"{name}_dt{dt}_o{order}_n"] = cnt
                out[f"{name}_dt{dt}_o{order}_pts"] = np.concatenate(P) if P else np.zeros((0, 3))
                out[f"{name}_dt{dt}_o{order}_wts"] = np.concatenate(W) if W else np.zeros(0)
    if only_return:
        return out
    np.savez_compressed(os.path.join(OUT, "quad_hex_rules.npz"), **out)
    print("quad_hex_rules.npz", len(out), "arrays")


if __name__ == "__main__":
    assert refload.available(), "build oracle/_ref first (bash oracle/build_ref.sh)"
    os.makedirs(OUT, exist_ok=True)
    if "--quad-hex-only" not in sys.argv:
        single_elements()
        small_meshes()
    quad_hex_rules()
