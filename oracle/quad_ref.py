"""TEST INFRASTRUCTURE -- not part of the product path.

SURVEY section 8 row f-5, groundwork: restatement of the reference's cut rule on the reference QUADRILATERAL
(LevelsetCutQuadrilateral, cutint/straightcutrule.cpp:206-501, policy FIND_OPTIMAL) and HEXAHEDRON for a multilinear level set
given by its vertex values (NGSolve vertex order (0,0), (1,0), (1,1), (0,1) [, the same at z = 1]).  Pure Python on top of the pinned 1D / 2D simplex
rules of oracle/ngsx_oracle.c: small cases only.  Pinned against tests/golden/quad_hex_rules.npz (rules produced by the compiled
reference, tools/gen_golden.py).  No CUDA path consumes this yet.
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


# ---------------------------------------------------------------------------------------------------------------------
# hexahedra (D = 3, xi = z): same algorithm one dimension up; the codimension-1 slices are the quadrilateral rule above
# ---------------------------------------------------------------------------------------------------------------------
def coeffs3(v):
    """LevelsetWrapper::GetCoeffsFromVals, ET_HEX (:513-525): lset = sum c[i][j][k] x^i y^j z^k"""
    c = np.zeros((2, 2, 2))
    c[0, 0, 0] = v[0]; c[1, 0, 0] = v[1] - v[0]; c[0, 1, 0] = v[3] - v[0]; c[0, 0, 1] = v[4] - v[0]
    c[1, 1, 0] = v[2] - c[1, 0, 0] - c[0, 1, 0] - c[0, 0, 0]
    c[1, 0, 1] = v[5] - c[1, 0, 0] - c[0, 0, 1] - c[0, 0, 0]
    c[0, 1, 1] = v[7] - c[0, 1, 0] - c[0, 0, 1] - c[0, 0, 0]
    c[1, 1, 1] = v[6] - c[1, 1, 0] - c[1, 0, 1] - c[0, 1, 1] - c[1, 0, 0] - c[0, 1, 0] - c[0, 0, 1] - c[0, 0, 0]
    return c


def lset3(c, x, y, z):
    """LevelsetWrapper::operator() (:538-542)"""
    v = 0.0
    for i in (0, 1):
        for j in (0, 1):
            for k in (0, 1):
                v += c[i, j, k] * x ** i * y ** j * z ** k
    return v


def grad3(c, x, y, z):
    """LevelsetWrapper::GetGrad (:527-533)"""
    return np.array([c[1, 0, 0] + c[1, 1, 0] * y + c[1, 0, 1] * z + c[1, 1, 1] * y * z,
                     c[0, 1, 0] + c[1, 1, 0] * x + c[0, 1, 1] * z + c[1, 1, 1] * x * z,
                     c[0, 0, 1] + c[0, 1, 1] * y + c[1, 0, 1] * x + c[1, 1, 1] * x * y])


def corners3(q):
    x0, x1, y0, y1, z0, z1 = q
    return [(x0, y0, z0), (x1, y0, z0), (x1, y1, z0), (x0, y1, z0), (x0, y0, z1), (x1, y0, z1), (x1, y1, z1), (x0, y1, z1)]


def hex_rule(order):
    p, w = segm_rule(order)
    n = len(p)
    pts = np.array([[p[i], p[j], p[k]] for i in range(n) for j in range(n) for k in range(n)]).reshape(-1, 3)
    return pts, np.array([w[i] * w[j] * w[k] for i in range(n) for j in range(n) for k in range(n)])


def plain_rule3(q, order):
    x0, x1, y0, y1, z0, z1 = q
    pts, w = hex_rule(order)
    vol = abs((x1 - x0) * (y1 - y0) * (z1 - z0))
    out = np.zeros_like(pts)
    out[:, 0] = x0 + (x1 - x0) * pts[:, 0]; out[:, 1] = y0 + (y1 - y0) * pts[:, 1]; out[:, 2] = z0 + (z1 - z0) * pts[:, 2]
    return list(out), list(w * vol)


def tensor_rule_along_xi3(c, dt, q, order):
    x0, x1, y0, y1, z0, z1 = q
    P, W = [], []
    ps, ws = segm_rule(order)
    for p1, w1 in zip(ps, ws):
        za = z0 + p1 * (z1 - z0)
        proj = [lset3(c, x0, y0, za), lset3(c, x1, y0, za), lset3(c, x1, y1, za), lset3(c, x0, y1, za)]
        p2, w2 = cut_quad_rule(proj, dt, order)                           # child LevelsetCutQuadrilateral on the unit square
        for pk, wk in zip(p2, w2):
            ip = np.array([x0 + pk[0] * (x1 - x0), y0 + pk[1] * (y1 - y0), za])
            scale = 1.0
            if dt == IF:
                g = grad3(c, *ip)
                scale = np.sqrt(g @ g) / np.sqrt(g[0] ** 2 + g[1] ** 2)
            P.append(ip); W.append(wk * w1 * (z1 - z0) * scale)
    return P, W


def rule_along_xi3(c, dt, q, order):
    cr = corners3(q)
    d = check([lset3(c, *p) for p in cr])
    P, W = [], []
    if d == IF:
        edges = ((0, 4), (1, 5), (2, 6), (3, 7))
        topo = any(check([lset3(c, *cr[b]), lset3(c, *cr[a])]) == IF for a, b in edges)
        if topo:
            xis = {0.0, 1.0}
            for a, b in edges:
                v = [lset3(c, *cr[b]), lset3(c, *cr[a])]
                if check(v) == IF:
                    xis.add(1.0 + v[0] / (v[0] - v[1]) * (0.0 - 1.0))
            xs = sorted(xis)
            for i in range(len(xs) - 1):
                sq = (q[0], q[1], q[2], q[3], xs[i], xs[i + 1])
                d2 = check([lset3(c, *p) for p in corners3(sq)], 1e-15)
                if d2 == IF:
                    p, w = tensor_rule_along_xi3(c, dt, sq, order)
                elif d2 == dt:
                    p, w = plain_rule3(sq, order)
                else:
                    p, w = [], []
                P += p; W += w
        else:
            P, W = tensor_rule_along_xi3(c, dt, q, order)
    elif d == dt:
        P, W = plain_rule3(q, order)
    return P, W


def sufficient_crit_bound(c):
    """GetSufficientCritsQBound (:353-390)"""
    cs = [(0, 0, 0), (1, 0, 0), (0, 1, 0), (1, 1, 0), (0, 0, 1), (1, 0, 1), (0, 1, 1), (1, 1, 1)]
    gs = [grad3(c, *p) for p in cs]
    vsq = sum(max(g[d] ** 2 for g in gs) for d in range(3))
    V = np.sqrt(vsq)
    with np.errstate(divide="ignore", invalid="ignore"):
        qmax = [0.0, 0.0, 0.0]
        for g in gs:
            for d in range(3):
                est = np.float64(V ** 2) / np.float64(V ** 2 - g[d] ** 2)
                if est > qmax[d]:
                    qmax[d] = est
        return [float(np.sqrt(1.0 - 1.0 / np.float64(x))) for x in qmax]


def tet_rule(tet, vals, dt, order):
    """LevelsetCutSimplex on a general tetrahedron with the P1 interpolant of `vals` (fallback path)"""
    Pn = [np.array(p, dtype=np.float64) for p in tet]
    det = abs(np.linalg.det(np.array([Pn[0] - Pn[3], Pn[1] - Pn[3], Pn[2] - Pn[3]])))
    if dt != IF:
        n, p, w = orc.cut_rule_untrafo(3, vals, dt, order)
        if n == -2:
            raise Throws("simplex")
        out = [p[k, 0] * Pn[0] + p[k, 1] * Pn[1] + p[k, 2] * Pn[2] + (1.0 - p[k, 0] - p[k, 1] - p[k, 2]) * Pn[3] for k in range(max(n, 0))]
        return out, [w[k] * det for k in range(max(n, 0))]
    cuts = []
    for i in range(4):
        for j in range(i + 1, 4):
            if (vals[i] >= 0) != (vals[j] >= 0):
                cuts.append(Pn[i] + vals[i] / (vals[i] - vals[j]) * (Pn[j] - Pn[i]))
    if len(cuts) == 3:
        trigs = [(cuts[0], cuts[1], cuts[2])]
    elif len(cuts) == 4:
        trigs = [(cuts[0], cuts[1], cuts[3]), (cuts[0], cuts[2], cuts[3])]   # Decompose (:131-196)
    else:
        raise Throws("simplex")
    rp = np.zeros((256, 3)); rw = np.zeros(256)
    n = orc.load().orc_select_rule(2, order, orc.P(rp), orc.P(rw), 256)
    out, ws = [], []
    for a, b, cc in trigs:
        fac = np.linalg.norm(np.cross(cc - a, b - a))                        # SimpleX::GetVolume (:66-72)
        for k in range(n):
            out.append((1.0 - rp[k, 0] - rp[k, 1]) * a + rp[k, 0] * b + rp[k, 1] * cc)
            ws.append(rw[k] * fac)
    return out, ws


def fallback_rule3(c, dt, q, order):
    cr = corners3(q)
    P, W = [], []
    for idx in ((3, 0, 1, 5), (3, 1, 2, 5), (3, 5, 2, 6), (4, 5, 0, 3), (4, 7, 5, 3), (7, 6, 5, 3)):
        tet = [cr[i] for i in idx]
        vals = []
        for p in tet:
            d = lset3(c, *p)
            vals.append(d if abs(d) > 1e-14 else 1e-14)
        d = check(vals)
        if d != IF and d == dt:
            Pn = [np.array(t, dtype=np.float64) for t in tet]
            det = abs(np.linalg.det(np.array([Pn[3] - Pn[0], Pn[2] - Pn[0], Pn[1] - Pn[0]])))
            rp = np.zeros((512, 3)); rw = np.zeros(512)
            n = orc.load().orc_select_rule(3, order, orc.P(rp), orc.P(rw), 512)
            p = [(1.0 - rp[k, 0] - rp[k, 1] - rp[k, 2]) * Pn[0] + rp[k, 0] * Pn[1] + rp[k, 1] * Pn[2] + rp[k, 2] * Pn[3] for k in range(n)]
            w = [rw[k] * det for k in range(n)]
        elif d == IF:
            p, w = tet_rule(tet, vals, dt, order)
        else:
            p, w = [], []
        P += p; W += w
    return P, W


def cut_hex_rule(vals, dt, order):
    """LevelsetCutQuadrilateral::GetIntegrationRule, D = 3, FIND_OPTIMAL (:436-451, :491-500)"""
    c = coeffs3(vals)
    q = (0.0, 1.0, 0.0, 1.0, 0.0, 1.0)
    sb = sufficient_crit_bound(c)
    if any(np.isnan(sb)):
        raise Throws("Sufficient Criterion calculated nan Bound!")
    md = 0
    for d in (1, 2):                                                       # min_element with a `<=` comparator: the last of equal minima
        if sb[d] <= sb[md]:
            md = d
    if not sb[md] < C_BOUND:
        return fallback_rule3(c, dt, q, order)
    if md == 2:
        return rule_along_xi3(c, dt, q, order)
    if md == 0:                                                            # X_Z (:322-336)
        P, W = rule_along_xi3(np.transpose(c, (2, 1, 0)).copy(), dt, q, order)
        return [np.array([p[2], p[1], p[0]]) for p in P], W
    P, W = rule_along_xi3(np.transpose(c, (0, 2, 1)).copy(), dt, q, order)  # Y_Z (:338-351)
    return [np.array([p[0], p[2], p[1]]) for p in P], W


def rule_untrafo_hex(vals, dt, order):
    """StraightCutIntegrationRuleUntransformed for ET_HEX: (n, points, weights); n = -1: no rule, -2: throws"""
    vals = [float(v) for v in vals]
    d = check(vals)
    try:
        if d == IF:
            P, W = cut_hex_rule(vals, dt, order)
            return len(W), np.array(P).reshape(-1, 3), np.array(W)
        if d != dt:
            return -1, np.zeros((0, 3)), np.zeros(0)
        p, w = hex_rule(order)
        return len(w), p, w
    except Throws:
        return -2, np.zeros((0, 3)), np.zeros(0)
