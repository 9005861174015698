"""Deterministic synthetic scenes restating BASELINE.json's five configs.

The reference's own scene files (e.g. the sandbox example config 1 names) are not mounted
(BASELINE.md §0), so every scene is generated: counter-based splitmix64 -> float, seed =
20261106 + config number, bodies spawned on a jittered grid with no initial overlap.

Each generator returns (bodies, joints) as numpy structured arrays of SfpBodyDesc / SfpJointDesc.
"""
from __future__ import annotations

import numpy as np

from ._ffi import (BODY_DESC_DTYPE, BODY_STATIC, JOINT_DESC_DTYPE, JOINT_DISTANCE, POLYGON_DTYPE,
                   SHAPE_CAPSULE, SHAPE_CIRCLE, SHAPE_POLYGON, SHAPE_RECT)

BASE_SEED = 20261106


def _splitmix64(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def uniform(seed: int, stream: int, n: int, lo: float = 0.0, hi: float = 1.0) -> np.ndarray:
    """n floats in [lo, hi): value i depends only on (seed, stream, i)."""
    with np.errstate(over="ignore"):
        ctr = np.arange(n, dtype=np.uint64) + np.uint64((stream & 0xFFFF) << 40)
        base = _splitmix64(np.full(1, seed, dtype=np.uint64))[0]
        bits = _splitmix64(ctr ^ base)
    u = (bits >> np.uint64(40)).astype(np.float64) / float(1 << 24)
    return (lo + (hi - lo) * u).astype(np.float32)


def empty_bodies(n: int) -> np.ndarray:
    b = np.zeros(n, dtype=BODY_DESC_DTYPE)
    b["density"] = 1.0
    b["mu_s"] = 0.5
    b["mu_d"] = 0.3
    b["restitution"] = 0.0
    b["category"] = 1
    b["mask"] = 0xFFFF
    return b


def static_rect(x, y, hx, hy, r=0.0, angle=0.0, mu_s=0.5, mu_d=0.3, e=0.0) -> np.ndarray:
    b = empty_bodies(1)
    b["x"], b["y"], b["angle"] = x, y, angle
    b["hx"], b["hy"], b["radius"] = hx, hy, r
    b["shape"] = SHAPE_RECT
    b["flags"] = BODY_STATIC
    b["mu_s"], b["mu_d"], b["restitution"] = mu_s, mu_d, e
    return b


def container(width: float, height: float, thickness: float = 0.5, **mat) -> np.ndarray:
    """floor + two walls, static rounded rects; interior is [-w/2, w/2] x [0, height]."""
    hw = 0.5 * width
    t = 0.5 * thickness
    return np.concatenate(
        [
            static_rect(0.0, -t, hw + thickness, t, 0.0, **mat),
            static_rect(-hw - t, 0.5 * height, t, 0.5 * height + thickness, 0.0, **mat),
            static_rect(hw + t, 0.5 * height, t, 0.5 * height + thickness, 0.0, **mat),
        ]
    )


def no_joints() -> np.ndarray:
    return np.zeros(0, dtype=JOINT_DESC_DTYPE)


def stack100(seed: int = BASE_SEED + 1):
    """Config 1: ground + 50 boxes (10 wide x 5 high brick stack) + 50 circles above."""
    ground = static_rect(0.0, -0.25, 10.0, 0.25, mu_s=0.6, mu_d=0.4, e=0.1)
    walls = np.concatenate([static_rect(-10.25, 3.0, 0.25, 3.5), static_rect(10.25, 3.0, 0.25, 3.5)])
    boxes = empty_bodies(50)
    ix = np.arange(50) % 10
    iy = np.arange(50) // 10
    boxes["shape"] = SHAPE_RECT
    boxes["hx"] = 0.23
    boxes["hy"] = 0.23
    boxes["radius"] = 0.02
    boxes["x"] = (ix - 4.5) * 0.55 + (iy % 2) * 0.12
    boxes["y"] = 0.25 + 0.002 + iy * 0.504
    circles = empty_bodies(50)
    circles["shape"] = SHAPE_CIRCLE
    circles["radius"] = 0.25
    circles["x"] = (ix - 4.5) * 0.6 + uniform(seed, 1, 50, -0.02, 0.02)
    circles["y"] = 0.25 + 5 * 0.504 + 1.0 + iy * 0.6
    for b in (boxes, circles):
        b["mu_s"], b["mu_d"], b["restitution"] = 0.6, 0.4, 0.1
    return np.concatenate([ground, boxes, circles, walls]), no_joints()


def circles_container(n: int = 10_000, seed: int = BASE_SEED + 2, width: float = 12.0):
    """Config 2: n circles r~U[0.05,0.10] falling into a static 3-piece container."""
    cols = int(round(np.sqrt(n)))
    rows = (n + cols - 1) // cols
    pitch = (width - 0.4) / cols
    r = uniform(seed, 1, n, 0.05, min(0.10, 0.45 * pitch))
    b = empty_bodies(n)
    i = np.arange(n)
    b["shape"] = SHAPE_CIRCLE
    b["radius"] = r
    b["x"] = (i % cols - 0.5 * (cols - 1)) * pitch + uniform(seed, 2, n, -0.1, 0.1) * pitch
    b["y"] = 0.3 + (i // cols) * max(pitch, 0.21) + uniform(seed, 3, n, 0.0, 0.05)
    b["mu_s"], b["mu_d"], b["restitution"] = 0.5, 0.3, 0.2
    height = 0.3 + rows * max(pitch, 0.21) + 1.0
    return np.concatenate([container(width, height, mu_s=0.5, mu_d=0.3, e=0.2), b]), no_joints()


def mixed_pile(n: int = 100_000, seed: int = BASE_SEED + 3, width: float | None = None,
               cols: int | None = None, rows_hint: int = 100):
    """Config 3: n bodies, type ~ uniform{circle, capsule, rounded rect}, random materials,
    spawned on a jittered grid (pitch 0.65 m > the largest bounding diameter, so nothing
    overlaps at spawn) about `rows_hint` rows high above a static floor with two walls.
    (SURVEY.md §8d suggested 400 x 250 rows; 250 rows make the top bodies fall 160 m and hit
    the pile at 55 m/s, which is a ballistics test, not a pile — 100 rows are used instead.)"""
    if cols is None:
        cols = max(1, int(np.ceil(n / rows_hint))) if n > rows_hint * rows_hint else max(
            1, int(round(np.sqrt(n * 1.6))))
    pitch = 0.65
    if width is None:
        width = cols * pitch + 0.6
    t = (uniform(seed, 1, n) * 3.0).astype(np.int32).clip(0, 2)
    b = empty_bodies(n)
    i = np.arange(n)
    circle, capsule, rect = t == 0, t == 1, t == 2
    b["shape"] = np.where(circle, SHAPE_CIRCLE, np.where(capsule, SHAPE_CAPSULE, SHAPE_RECT))
    b["radius"] = np.where(
        circle, uniform(seed, 2, n, 0.05, 0.12),
        np.where(capsule, uniform(seed, 3, n, 0.03, 0.08), uniform(seed, 4, n, 0.01, 0.04)))
    b["hx"] = np.where(circle, 0.0, uniform(seed, 5, n, 0.05, 0.20))
    b["hy"] = np.where(rect, uniform(seed, 6, n, 0.05, 0.20), 0.0)
    b["angle"] = uniform(seed, 7, n, 0.0, 2.0 * np.pi)
    b["mu_s"] = uniform(seed, 8, n, 0.4, 0.8)
    b["mu_d"] = 0.7 * b["mu_s"]
    b["restitution"] = uniform(seed, 9, n, 0.0, 0.4)
    b["x"] = (i % cols - 0.5 * (cols - 1)) * pitch + uniform(seed, 10, n, -0.02, 0.02)
    b["y"] = 0.35 + (i // cols) * pitch + uniform(seed, 11, n, 0.0, 0.02)
    rows = (n + cols - 1) // cols
    height = 0.35 + rows * pitch + 1.0
    return np.concatenate([container(width, height, mu_s=0.6, mu_d=0.42, e=0.1), b]), no_joints()


def shelves(n: int = 4_000_000, levels: int = 64, rows: int = 32, seed: int = BASE_SEED + 6):
    """Scale point beyond BASELINE.json's configs: n mixed bodies spread
    over `levels` static floors stacked 24 m apart, `rows` spawn rows per floor, so the whole
    scene stays within ~1.5 km of the origin (FP32 positions) while its state (n x 88 B plus
    pair data) is several times the 126 MB L2 — the size at which HBM bandwidth, not phase
    latency, bounds the step."""
    per_level = (n + levels - 1) // levels
    cols = (per_level + rows - 1) // rows
    pitch, gap = 0.65, 24.0
    width = cols * pitch + 0.6
    parts = []
    for lv in range(levels):
        parts.append(static_rect(0.0, lv * gap - 0.25, 0.5 * width + 0.5, 0.25, mu_s=0.6, mu_d=0.42,
                                 e=0.1))
    parts.append(static_rect(-0.5 * width - 0.25, 0.5 * levels * gap, 0.25, 0.5 * levels * gap + 1.0))
    parts.append(static_rect(0.5 * width + 0.25, 0.5 * levels * gap, 0.25, 0.5 * levels * gap + 1.0))
    b, _ = mixed_pile(n, seed=seed, cols=cols, width=width)
    b = b[3:]
    i = np.arange(n)
    lv = i // per_level
    k = i % per_level
    b["x"] = (k % cols - 0.5 * (cols - 1)) * pitch + uniform(seed, 20, n, -0.02, 0.02)
    b["y"] = lv * gap + 0.35 + (k // cols) * pitch + uniform(seed, 21, n, 0.0, 0.02)
    parts.append(b)
    return np.concatenate(parts), no_joints()


def ropes(n_ropes: int = 2000, links: int = 64, n_free: int = 20_000, seed: int = BASE_SEED + 4,
          rope_rows: int = 4):
    """Config 4: n_ropes chains of `links` particles (circle r=0.03, spacing 0.08) hanging from
    static anchors (rope_rows tiers of anchors), joined by distance constraints, plus free
    circles spawned in the gaps between the ropes over the whole height, and a floor.
    (SURVEY.md §8d put the anchors on a 50 x 40 grid, i.e. 40 tiers = 230 m of height: the
    free circles then arrive at 65 m/s.  4 tiers keep every speed below ~25 m/s.)"""
    spacing = 0.08
    arows = max(1, min(rope_rows, n_ropes))
    acols = (n_ropes + arows - 1) // arows
    ax_pitch = 0.5
    rope_len = links * spacing
    row_pitch = rope_len + 0.6
    width = acols * ax_pitch + 2.0
    top = arows * row_pitch + 1.0
    parts = [container(width, top + 4.0)]
    n_static0 = 3
    # anchors: small static circles on a layer that collides with nothing
    anchors = empty_bodies(n_ropes)
    r_i = np.arange(n_ropes)
    anchors["shape"] = SHAPE_CIRCLE
    anchors["radius"] = 0.02
    anchors["x"] = (r_i % acols - 0.5 * (acols - 1)) * ax_pitch
    anchors["y"] = top - (r_i // acols) * row_pitch
    anchors["flags"] = BODY_STATIC
    anchors["category"] = 2
    anchors["mask"] = 0
    parts.append(anchors)
    # links hang straight down from each anchor
    nl = n_ropes * links
    lk = empty_bodies(nl)
    rope_of = np.arange(nl) // links
    k = np.arange(nl) % links
    lk["shape"] = SHAPE_CIRCLE
    lk["radius"] = 0.03
    lk["x"] = anchors["x"][rope_of] + uniform(seed, 1, nl, -0.001, 0.001)
    lk["y"] = anchors["y"][rope_of] - (k + 1) * spacing
    lk["mu_s"], lk["mu_d"] = 0.4, 0.3
    parts.append(lk)
    # free circles in the gaps between rope columns, spread over the height
    fr = empty_bodies(n_free)
    fi = np.arange(n_free)
    frows = max(1, (n_free + acols - 1) // acols)
    fr["shape"] = SHAPE_CIRCLE
    fr["radius"] = uniform(seed, 2, n_free, 0.05, 0.10)
    fr["x"] = (fi % acols - 0.5 * (acols - 1)) * ax_pitch + 0.25 + uniform(seed, 3, n_free, -0.05, 0.05)
    fr["y"] = 0.5 + (fi // acols) * ((top + 2.5) / frows) + uniform(seed, 4, n_free, 0.0, 0.02)
    parts.append(fr)
    bodies = np.concatenate(parts)
    link0 = n_static0 + n_ropes
    j = np.zeros(nl, dtype=JOINT_DESC_DTYPE)
    j["kind"] = JOINT_DISTANCE
    j["rest"] = spacing
    j["compliance"] = 0.0
    body = link0 + np.arange(nl)
    j["body_b"] = body
    j["body_a"] = np.where(k == 0, n_static0 + rope_of, body - 1)
    return bodies, j


def small_world(world: int, n_bodies: int = 512, seed: int = BASE_SEED + 5,
                width: float = 16.0):
    """Config 5, one world: config-3 distribution scaled down; 3 static + (n-3) dynamic."""
    nd = n_bodies - 3
    cols = max(1, int((width - 0.6) / 0.65))
    b, _ = mixed_pile(nd, seed=(seed * 1_000_003 + world) & 0x7FFFFFFFFFFF, width=width, cols=cols)
    return b, no_joints()


def worlds(n_worlds: int, n_bodies: int = 512, first_world: int = 0, seed: int = BASE_SEED + 5):
    """Config 5: world-major array (n_worlds, n_bodies) of body descs."""
    out = np.zeros((n_worlds, n_bodies), dtype=BODY_DESC_DTYPE)
    for w in range(n_worlds):
        out[w] = small_world(first_world + w, n_bodies, seed)[0]
    return out


def random_soup(n: int, seed: int, pitch: float = 0.7, extent: float | None = None):
    """Unstructured test scene: mixed shapes on a jittered grid (neighbours overlap slightly),
    random poses and velocities, ~10% static bodies, several collision layers, a floor — for
    parity and property tests.  `extent` (if given) scatters uniformly in a box instead."""
    b = empty_bodies(n)
    t = (uniform(seed, 1, n) * 3.0).astype(np.int32).clip(0, 2)
    b["shape"] = t
    b["radius"] = np.where(t == 2, uniform(seed, 2, n, 0.0, 0.05), uniform(seed, 2, n, 0.04, 0.12))
    b["hx"] = np.where(t == 0, 0.0, uniform(seed, 3, n, 0.05, 0.3))
    b["hy"] = np.where(t == 2, uniform(seed, 4, n, 0.05, 0.25), 0.0)
    cols = max(1, int(np.ceil(np.sqrt(n))))
    i = np.arange(n)
    if extent is None:
        b["x"] = (i % cols - 0.5 * (cols - 1)) * pitch + uniform(seed, 5, n, -0.1, 0.1)
        b["y"] = 0.5 + (i // cols) * pitch + uniform(seed, 6, n, -0.1, 0.1)
    else:
        b["x"] = uniform(seed, 5, n, -extent, extent)
        b["y"] = uniform(seed, 6, n, 0.5, 0.5 + extent)
    b["angle"] = uniform(seed, 7, n, -3.0, 3.0)
    b["vx"] = uniform(seed, 8, n, -1.0, 1.0)
    b["vy"] = uniform(seed, 9, n, -1.0, 1.0)
    b["omega"] = uniform(seed, 10, n, -3.0, 3.0)
    b["mu_s"] = uniform(seed, 11, n, 0.0, 0.9)
    b["mu_d"] = 0.7 * b["mu_s"]
    b["restitution"] = uniform(seed, 12, n, 0.0, 0.6)
    st = uniform(seed, 13, n) < 0.1
    b["flags"] = np.where(st, BODY_STATIC, 0)
    lay = (uniform(seed, 14, n) * 4.0).astype(np.int32).clip(0, 3)
    b["category"] = (1 << lay).astype(np.uint16)
    b["mask"] = np.where(lay == 3, 0x7, 0xFFFF).astype(np.uint16)
    floor = static_rect(0.0, -0.5, 0.5 * cols * pitch + 2.0, 0.5)
    return np.concatenate([b, floor]), no_joints()


def polygon(points) -> np.ndarray:
    """One SfpPolygon from CCW points, re-centred on its centroid (as the ABI requires)."""
    pts = np.asarray(points, dtype=np.float64)
    x, y = pts[:, 0], pts[:, 1]
    xn, yn = np.roll(x, -1), np.roll(y, -1)
    cr = x * yn - xn * y
    a2 = cr.sum()
    cx, cy = ((x + xn) * cr).sum() / (3 * a2), ((y + yn) * cr).sum() / (3 * a2)
    p = np.zeros(1, dtype=POLYGON_DTYPE)
    p["count"] = len(pts)
    p["x"][0, : len(pts)] = x - cx
    p["y"][0, : len(pts)] = y - cy
    return p


def regular_polygon(n_verts: int, radius: float, phase: float = 0.0) -> np.ndarray:
    a = phase + 2 * np.pi * np.arange(n_verts) / n_verts
    return polygon(np.stack([radius * np.cos(a), radius * np.sin(a)], axis=1))


def polygon_table(seed: int = BASE_SEED + 8) -> np.ndarray:
    """a small palette: triangle, irregular quad, pentagon, hexagon, heptagon, octagon, sliver"""
    return np.concatenate([
        regular_polygon(3, 0.18), polygon([(-0.2, -0.1), (0.15, -0.14), (0.22, 0.08), (-0.1, 0.15)]),
        regular_polygon(5, 0.16, 0.3), regular_polygon(6, 0.15), regular_polygon(7, 0.17, 0.1),
        regular_polygon(8, 0.14), polygon([(-0.3, -0.04), (0.3, -0.04), (0.3, 0.04), (-0.3, 0.04)]),
    ])


def polygon_pile(n: int = 1500, seed: int = BASE_SEED + 8, cols: int | None = None):
    """mixed pile in which ~half of the bodies are general convex polygons (with and without
    rounding).  Returns (bodies, joints, polygons)."""
    table = polygon_table(seed)
    b, j = mixed_pile(n, seed=seed, cols=cols)
    k = np.arange(len(b)) >= 3
    pick = uniform(seed, 30, len(b)) < 0.5
    which = (uniform(seed, 31, len(b)) * len(table)).astype(np.uint32).clip(0, len(table) - 1)
    sel = k & pick
    b["shape"] = np.where(sel, SHAPE_POLYGON | (which << 8), b["shape"])
    b["radius"] = np.where(sel, np.where(uniform(seed, 32, len(b)) < 0.5, 0.0, 0.02), b["radius"])
    b["hx"] = np.where(sel, 0.0, b["hx"])
    b["hy"] = np.where(sel, 0.0, b["hy"])
    return b, j, table


def crowded_plank(n_riders: int = 100, seed: int = BASE_SEED + 7):
    """One long DYNAMIC plank resting on a floor with n_riders small bodies standing on it: the
    plank has more partners than there are colours (63), which exercises the serial group."""
    pitch = 0.13
    hx = 0.5 * n_riders * pitch + 0.2
    floor = static_rect(0.0, -0.5, hx + 2.0, 0.5)
    plank = empty_bodies(1)
    plank["shape"], plank["hx"], plank["hy"], plank["radius"] = SHAPE_RECT, hx, 0.05, 0.01
    plank["x"], plank["y"] = 0.0, 0.062
    riders = empty_bodies(n_riders)
    k = np.arange(n_riders)
    riders["shape"] = np.where(k % 3 == 0, SHAPE_RECT, SHAPE_CIRCLE)
    riders["radius"] = np.where(k % 3 == 0, 0.01, 0.05)
    riders["hx"] = np.where(k % 3 == 0, 0.04, 0.0)
    riders["hy"] = np.where(k % 3 == 0, 0.04, 0.0)
    riders["x"] = (k - 0.5 * (n_riders - 1)) * pitch + uniform(seed, 1, n_riders, -0.005, 0.005)
    riders["y"] = 0.062 + 0.06 + 0.052 + uniform(seed, 2, n_riders, 0.0, 0.01)
    return np.concatenate([floor, plank, riders]), no_joints()


CONFIGS = {
    "stack100": stack100,
    "circles10k": circles_container,
    "pile100k": mixed_pile,
    "ropes2k": ropes,
}
