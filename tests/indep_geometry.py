"""Float64 geometry and solver formulas written WITHOUT include/sfp_math.h, for the independent
checks of tests/test_independent.py.

The oracle and the CUDA kernels compile the same header, so "GPU == oracle bit for bit" cannot
catch a formula that is wrong on both sides.  What is here shares nothing with that header: no
SAT, no clipping, no reference face.  Separation and normal of two rounded convex shapes come
from the exact signed distance function of one shape evaluated on a dense sampling of the
other's boundary (and the other way round); the position solve is the XPBD contact update of
Mueller et al. 2020 (section 3.3.1 / 3.5) written out for 2D with numpy.
"""
from __future__ import annotations

import numpy as np

CIRCLE, CAPSULE, RECT, POLYGON = 0, 1, 2, 3


def core_vertices(kind: int, hx: float, hy: float, poly=None) -> np.ndarray:
    """body-frame vertices of the collider's core, counter-clockwise"""
    if kind == CIRCLE:
        return np.zeros((1, 2))
    if kind == CAPSULE:
        return np.array([[-hx, 0.0], [hx, 0.0]])
    if kind == RECT:
        return np.array([[-hx, -hy], [hx, -hy], [hx, hy], [-hx, hy]])
    n = int(poly["count"])
    return np.stack([poly["x"][:n].astype(np.float64), poly["y"][:n].astype(np.float64)], axis=1)


def to_world(verts: np.ndarray, x: float, y: float, angle: float) -> np.ndarray:
    c, s = np.cos(angle), np.sin(angle)
    return np.stack([x + c * verts[:, 0] - s * verts[:, 1], y + s * verts[:, 0] + c * verts[:, 1]], axis=1)


def sdf_core(P: np.ndarray, V: np.ndarray):
    """exact signed distance of the points P (m, 2) to the convex core with world vertices V, and
    its gradient (unit vector pointing away from the core)"""
    n = len(V)
    if n == 1:
        d = P - V[0]
        dist = np.hypot(d[:, 0], d[:, 1])
        g = d / np.maximum(dist, 1e-300)[:, None]
        return dist, g
    A = V
    B = np.roll(V, -1, axis=0)
    if n == 2:
        A, B = V[:1], V[1:]
    E = B - A                                   # (e, 2)
    ee = (E * E).sum(1)
    W = P[:, None, :] - A[None, :, :]           # (m, e, 2)
    u = np.clip((W * E[None]).sum(2) / ee[None], 0.0, 1.0)
    C = A[None] + u[..., None] * E[None]        # closest point on every edge
    D = P[:, None, :] - C
    dist = np.hypot(D[..., 0], D[..., 1])       # (m, e)
    k = dist.argmin(1)
    rows = np.arange(len(P))
    dmin = dist[rows, k]
    g = D[rows, k] / np.maximum(dmin, 1e-300)[:, None]
    if n == 2:
        return dmin, g
    # inside test: left of every edge of the CCW polygon
    N = np.stack([E[:, 1], -E[:, 0]], axis=1) / np.sqrt(ee)[:, None]   # outward normals
    S = (W * N[None]).sum(2)                    # signed distance to every edge line
    inside = (S <= 0.0).all(1)
    kin = S.argmax(1)
    sd = np.where(inside, S[rows, kin], dmin)
    g = np.where(inside[:, None], N[kin], g)
    return sd, g


def boundary_samples(V: np.ndarray, r: float, per_edge: int = 192, per_arc: int = 96) -> np.ndarray:
    """points on the boundary of the core V inflated by r"""
    n = len(V)
    if n == 1:
        a = np.linspace(0.0, 2 * np.pi, 4 * per_arc, endpoint=False)
        return V[0] + r * np.stack([np.cos(a), np.sin(a)], axis=1)
    idx = [(0, 1), (1, 0)] if n == 2 else [(i, (i + 1) % n) for i in range(n)]
    t = np.linspace(0.0, 1.0, per_edge)
    pts = []
    normals = []
    for i, j in idx:
        e = V[j] - V[i]
        nrm = np.array([e[1], -e[0]]) / np.hypot(*e)
        normals.append(nrm)
        pts.append(V[i] + t[:, None] * e + r * nrm)
    if r > 0.0:
        for q, (i, j) in enumerate(idx):
            # arc around vertex V[j] from this edge's normal to the next edge's normal
            n0, n1 = normals[q], normals[(q + 1) % len(idx)]
            a0 = np.arctan2(n0[1], n0[0])
            da = (np.arctan2(n1[1], n1[0]) - a0) % (2 * np.pi)
            a = a0 + np.linspace(0.0, da, per_arc)
            pts.append(V[j] + r * np.stack([np.cos(a), np.sin(a)], axis=1))
    return np.concatenate(pts)


def separation(VA, ra, VB, rb, directions: int = 3600, **_):
    """(signed separation, unit normal from A to B) of two rounded convex shapes from their support
    functions over a dense fan of directions, refined twice around the best one:
        sep = max over unit n of  min_{b in B} b.n - max_{a in A} a.n.
    Positive: the distance between the shapes.  Negative: minus the smallest translation that
    separates them (the penetration depth a contact solver has to remove)."""
    lo, hi = 0.0, 2.0 * np.pi
    best, best_a = -np.inf, 0.0
    for _level in range(3):
        a = np.linspace(lo, hi, directions, endpoint=False)
        N = np.stack([np.cos(a), np.sin(a)], axis=1)
        sep = (VB @ N.T).min(0) - (VA @ N.T).max(0) - (ra + rb)
        k = int(sep.argmax())
        if sep[k] > best:
            best, best_a = float(sep[k]), float(a[k])
        step = (hi - lo) / directions
        lo, hi = best_a - 2.0 * step, best_a + 2.0 * step
    return best, np.array([np.cos(best_a), np.sin(best_a)])


def separation_sdf(VA, ra, VB, rb, **kw):
    """the same for DISJOINT shapes by a different route (signed distance function of one shape
    on a dense sampling of the other's boundary, both ways); for overlapping shapes it measures
    the depth of the deepest boundary point instead, which is smaller in corner-in-corner
    configurations"""
    pb = boundary_samples(VB, rb, **kw)
    sa, ga = sdf_core(pb, VA)
    ia = sa.argmin()
    pa = boundary_samples(VA, ra, **kw)
    sb, gb = sdf_core(pa, VB)
    ib = sb.argmin()
    s1, s2 = sa[ia] - ra, sb[ib] - rb
    if s1 <= s2:
        return s1, ga[ia]
    return s2, -gb[ib]


# ---- mass properties (textbook solids; the rounded rectangle as its bounding rectangle, which is
# the library's stated approximation) -------------------------------------------------------------
def inverse_mass(kind, hx, hy, r, density):
    if kind == CIRCLE:
        m = density * np.pi * r * r
        inertia = 0.5 * m * r * r
    elif kind == CAPSULE:
        length = 2.0 * hx
        mb, mc = density * length * 2.0 * r, density * np.pi * r * r
        lc = 4.0 * r / (3.0 * np.pi)
        m = mb + mc
        inertia = mb * (length ** 2 + 4.0 * r * r) / 12.0 + mc * (0.5 * r * r + hx * hx + 2.0 * hx * lc)
    else:
        w, h = 2.0 * (hx + r), 2.0 * (hy + r)
        m = density * w * h
        inertia = m * (w * w + h * h) / 12.0
    return 1.0 / m, 1.0 / inertia


def cross(a, b):
    return a[0] * b[1] - a[1] * b[0]


def xpbd_contact_1pt(xa, tha, ima, iia, xb, thb, imb, iib, ra, rb, n, d):
    """One XPBD non-penetration projection (compliance 0) for a contact with lever arms ra, rb
    (contact point minus centre of mass), unit normal n from A to B and penetration depth d.
    Returns the new (xa, tha, xb, thb)."""
    if d <= 0.0:
        return xa, tha, xb, thb
    wa = ima + iia * cross(ra, n) ** 2
    wb = imb + iib * cross(rb, n) ** 2
    dl = d / (wa + wb)
    p = dl * n
    return xa - ima * p, tha - iia * cross(ra, p), xb + imb * p, thb + iib * cross(rb, p)


def xpbd_contact_2pt(xa, tha, ima, iia, xb, thb, imb, iib, ras, rbs, n, d):
    """Both points of a face contact solved together (the library's stated choice: a 2x2 system
    with both multipliers non-negative); None when the block solve is not the active set, which
    the caller then skips."""
    d = np.asarray(d, dtype=np.float64)
    ka = np.array([cross(r, n) for r in ras])
    kb = np.array([cross(r, n) for r in rbs])
    K = (ima + imb) + iia * np.outer(ka, ka) + iib * np.outer(kb, kb)
    if np.linalg.det(K) <= 1e-2 * K[0, 0] * K[1, 1]:
        return None
    lam = np.linalg.solve(K, d)
    if (lam < 0.0).any():
        return None
    p = lam.sum() * n
    return xa - ima * p, tha - iia * (ka @ lam), xb + imb * p, thb + iib * (kb @ lam)
