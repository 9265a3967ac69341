"""TEST INFRASTRUCTURE -- not part of the product path.

SURVEY section 8 row f-5, groundwork: restatement of the reference's cut rule on the reference QUADRILATERAL
(LevelsetCutQuadrilateral, cutint/straightcutrule.cpp:206-501, policy FIND_OPTIMAL, 2D only) for a bilinear level set given
by its four vertex values (NGSolve vertex order (0,0), (1,0), (1,1), (0,1)).  Pure Python on top of the pinned 1D / 2D simplex
rules of oracle/ngsx_oracle.c: small cases only.  Pinned against tests/golden/quad_hex_rules.npz (rules produced by the compiled
reference, tools/gen_golden.py).  No CUDA path consumes this yet; hexahedra are not restated.
"""
import numpy as np

from . import orc

NEG, POS, IF = 0, 1, 2
C_BOUND = 0.999                                  # straightcutrule.cpp:252 `c`


class Throws(Exception):
    """the reference throws on this input"""


def check(vals, eps=0.0):
    """CheckIfStraightCut, cutint/straightcutrule.cpp:11-39"""
    haspos = any(v > eps for v in vals)
    hasneg = any(v < -eps for v in vals)
    if haspos == hasneg:
        return IF
    return NEG if hasneg else POS


def coeffs(v):
    """LevelsetWrapper::GetCoeffsFromVals, ET_QUAD (:503-520): lset = c00 + c10 x + c01 y + c11 x y"""
    c00 = v[0]; c10 = v[1] - c00; c01 = v[3] - c00; c11 = v[2] - c10 - c01 - c00
    return np.array([c00, c10, c01, c11], dtype=np.float64)


def lset(c, x, y):
    return c[0] + c[1] * x + c[2] * y + c[3] * x * y


def grad(c, x, y):
    return np.array([c[1] + c[3] * y, c[2] + c[3] * x])


def segm_rule(order):
    pts = np.zeros((64, 3)); w = np.zeros(64)
    n = orc.load().orc_select_rule(1, order, orc.P(pts), orc.P(w), 64)
    return pts[:n, 0].copy(), w[:n].copy()


def quad_rule(order):
    """SelectIntegrationRule(ET_QUAD, order): tensor Gauss rule, x outer, y inner (checked against the compiled reference)"""
    p, w = segm_rule(order)
    pts = np.array([[p[i], p[j], 0.0] for i in range(len(p)) for j in range(len(p))]).reshape(-1, 3)
    return pts, np.array([w[i] * w[j] for i in range(len(p)) for j in range(len(p))])


def corners(q):
    x0, x1, y0, y1 = q
    return [(x0, y0), (x1, y0), (x1, y1), (x0, y1)]                      # Quadrilateral(bnds), straightcutrule.hpp:71-77


def plain_rule(q, order):
    """Quadrilateral::GetPlainIntegrationRule (:102-128)"""
    x0, x1, y0, y1 = q
    pts, w = quad_rule(order)
    vol = abs((x1 - x0) * (y1 - y0))
    out = np.zeros_like(pts)
    out[:, 0] = x0 + (x1 - x0) * pts[:, 0]
    out[:, 1] = y0 + (y1 - y0) * pts[:, 1]
    return list(out), list(w * vol)


def tensor_rule_along_xi(c, dt, q, order):
    """GetTensorProductAlongXiIntegrationRule (:255-303), D = 2 (xi = y)"""
    x0, x1, y0, y1 = q
    P, W = [], []
    ps, ws = segm_rule(order)
    for p1, w1 in zip(ps, ws):
        ya = y0 + p1 * (y1 - y0)
        proj = [lset(c, x1, ya), lset(c, x0, ya)]                          # ET_SEGM vertex 0 sits at xhat = 1
        if check(proj) != IF or 0.0 in proj:
            raise Throws("LevelsetCutSimplex::Decompose on an uncut segment")
        n, p2, w2 = orc.cut_rule_untrafo(1, proj, dt, order)
        for k in range(max(n, 0)):
            x = x0 + p2[k, 0] * (x1 - x0)
            scale = 1.0
            if dt == IF:
                g = grad(c, x, ya)
                scale = np.sqrt(g[0] ** 2 + g[1] ** 2) / abs(g[0])
            P.append(np.array([x, ya, 0.0])); W.append(w2[k] * w1 * (y1 - y0) * scale)
    return P, W


def has_topology_change(c, q):
    cr = corners(q)
    for a, b in ((1, 2), (0, 3)):                                          # :206-218
        if check([lset(c, *cr[b]), lset(c, *cr[a])]) == IF:
            return True
    return False


def decompose(c, q):
    """LevelsetCutQuadrilateral::Decompose (:220-251): cut the quad at the heights where the level set changes sign on a vertical edge"""
    cr = corners(q)
    xis = {0.0, 1.0}
    for a, b in ((1, 2), (0, 3)):
        v = [lset(c, *cr[b]), lset(c, *cr[a])]
        if check(v) == IF:
            xis.add(1.0 + v[0] / (v[0] - v[1]) * (0.0 - 1.0))              # CalcIFPolytopEUsingLset on the unit line (points 1, 0)
    xs = sorted(xis)
    return [(q[0], q[1], xs[i], xs[i + 1]) for i in range(len(xs) - 1)]


def rule_along_xi(c, dt, q, order):
    """GetIntegrationRuleAlongXi (:453-468)"""
    d = check([lset(c, *p) for p in corners(q)])
    P, W = [], []
    if d == IF:
        if has_topology_change(c, q):
            for sq in decompose(c, q):
                d2 = check([lset(c, *p) for p in corners(sq)], 1e-15)
                if d2 == IF:
                    p, w = tensor_rule_along_xi(c, dt, sq, order)
                elif d2 == dt:
                    p, w = plain_rule(sq, order)
                else:
                    p, w = [], []
                P += p; W += w
        else:
            P, W = tensor_rule_along_xi(c, dt, q, order)
    elif d == dt:
        P, W = plain_rule(q, order)
    return P, W


def exact_crit_bound_2d(c):
    """GetExactCritsQBound2D (:392-413)"""
    allow = [True, True]
    with np.errstate(divide="ignore", invalid="ignore"):
        h = -np.float64(c[1]) / np.float64(c[3])
        if 0 < h < 1:
            allow[1] = False
        h = -np.float64(c[2]) / np.float64(c[3])
        if 0 < h < 1:
            allow[0] = False
        qmax = [0.0, 0.0]
        for p in ((0, 0), (1, 0), (1, 1), (0, 1)):
            g = grad(c, *p)
            nrm = np.sqrt(g[0] ** 2 + g[1] ** 2)
            qy, qx = abs(g[1]) / nrm, abs(g[0]) / nrm
            if qy > qmax[1]:
                qmax[1] = qy
            if qx > qmax[0]:
                qmax[0] = qx
    for k in (0, 1):
        if not allow[k]:
            qmax[k] = 2.0
    return qmax


def triangle_rule(tri, vals, dt, order):
    """LevelsetCutSimplex on a general triangle whose level set is the P1 interpolant of `vals` (fallback path): volumes by the
    affine image of the pinned reference-triangle rule; interface segments with their length in the ambient coordinates"""
    P0, P1, P2 = (np.array(p, dtype=np.float64) for p in tri)
    det = abs((P0[0] - P2[0]) * (P1[1] - P2[1]) - (P1[0] - P2[0]) * (P0[1] - P2[1]))
    if dt != IF:
        n, p, w = orc.cut_rule_untrafo(2, vals, dt, order)
        if n == -2:
            raise Throws("simplex")
        out = []
        for k in range(max(n, 0)):
            x = p[k, 0] * P0 + p[k, 1] * P1 + (1.0 - p[k, 0] - p[k, 1]) * P2
            out.append(np.array([x[0], x[1], 0.0]))
        return out, [w[k] * det for k in range(max(n, 0))]
    # interface: the two zeros of the P1 level set on the edges (pairs i < j with a sign change, :41-64), one segment
    pts = [P0, P1, P2]
    cuts = []
    for i in range(3):
        for j in range(i + 1, 3):
            if (vals[i] >= 0) != (vals[j] >= 0):
                cuts.append(pts[i] + vals[i] / (vals[i] - vals[j]) * (pts[j] - pts[i]))
    if len(cuts) != 2:
        raise Throws("simplex")
    ps, ws = segm_rule(order)
    length = np.linalg.norm(cuts[1] - cuts[0])
    out = [np.array([*((1.0 - s) * cuts[0] + s * cuts[1]), 0.0]) for s in ps]
    return out, [wq * length for wq in ws]


def fallback_rule(c, dt, q, order):
    """GetFallbackIntegrationRule (:470-489), D = 2: two triangles, level set re-interpolated at their vertices (clamped at 1e-14)"""
    cr = corners(q)
    P, W = [], []
    for idx in ((0, 1, 3), (2, 1, 3)):
        tri = [cr[i] for i in idx]
        vals = []
        for p in tri:
            d = lset(c, *p)
            vals.append(d if abs(d) > 1e-14 else 1e-14)                    # LevelsetWrapper::update_initial_coefs (:547-556)
        d = check(vals)
        if d != IF and d == dt:                                            # SimpleX::GetPlainIntegrationRule (:80-100): base point p0
            P0, P1, P2 = (np.array(t, dtype=np.float64) for t in tri)
            det = abs((P1[0] - P0[0]) * (P2[1] - P0[1]) - (P2[0] - P0[0]) * (P1[1] - P0[1]))
            rp = np.zeros((256, 3)); rw = np.zeros(256)
            n = orc.load().orc_select_rule(2, order, orc.P(rp), orc.P(rw), 256)
            p = [np.array([*((1.0 - rp[k, 0] - rp[k, 1]) * P0 + rp[k, 0] * P1 + rp[k, 1] * P2), 0.0]) for k in range(n)]
            w = [rw[k] * det for k in range(n)]
        elif d == IF:
            p, w = triangle_rule(tri, vals, dt, order)
        else:
            p, w = [], []
        P += p; W += w
    return P, W


def cut_quad_rule(vals, dt, order):
    """LevelsetCutQuadrilateral::GetIntegrationRule (:491-500) on the reference square, policy FIND_OPTIMAL (:415-434)"""
    c = coeffs(vals)
    q = (0.0, 1.0, 0.0, 1.0)
    e = exact_crit_bound_2d(c)
    if e[0] < C_BOUND and e[1] < C_BOUND:
        sw = "ID" if e[1] <= e[0] else "XY"
    elif e[1] < C_BOUND:
        sw = "ID"
    elif e[0] < C_BOUND:
        sw = "XY"
    else:
        sw = "NONE"
    if sw == "ID":
        return rule_along_xi(c, dt, q, order)
    if sw == "XY":                                                        # GetIntegrationRuleOnXYPermutatedQuad (:305-320)
        cr = np.array([c[0], c[2], c[1], c[3]])
        P, W = rule_along_xi(cr, dt, q, order)
        return [np.array([p[1], p[0], p[2]]) for p in P], W
    return fallback_rule(c, dt, q, order)


def rule_untrafo(vals, dt, order):
    """StraightCutIntegrationRuleUntransformed for ET_QUAD (:663-726): (n, points [n][3], weights [n]); n = -1: no rule, -2: throws"""
    vals = [float(v) for v in vals]
    d = check(vals)
    try:
        if d == IF:
            P, W = cut_quad_rule(vals, dt, order)
            return len(W), np.array(P).reshape(-1, 3), np.array(W)
        if d != dt:
            return -1, np.zeros((0, 3)), np.zeros(0)
        p, w = quad_rule(order)
        return len(w), p, w
    except Throws:
        return -2, np.zeros((0, 3)), np.zeros(0)
