"""Spatial queries: ray casts, AABB, shape and point queries.  CPU tier: closed-form
answers on the oracle.  GPU tier: the device's hash walk returns exactly what the oracle's brute
force over all bodies returns."""
import math

import numpy as np
import pytest

import starframe_b200 as sf
from starframe_b200 import _ffi, scenes
from starframe_b200._ffi import RAY_DTYPE, SHAPE_CAPSULE, SHAPE_CIRCLE, SHAPE_RECT

NONE = 0xFFFFFFFF


def body(shape, x, y, angle=0.0, hx=0.0, hy=0.0, r=0.0, **kw):
    b = scenes.empty_bodies(1)
    b["shape"], b["x"], b["y"], b["angle"] = shape, x, y, angle
    b["hx"], b["hy"], b["radius"] = hx, hy, r
    for k, v in kw.items():
        b[k] = v
    return b


def ray(ox, oy, dx, dy, max_t=100.0, mask=0xFFFF):
    r = np.zeros(1, dtype=RAY_DTYPE)
    r["ox"], r["oy"], r["dx"], r["dy"], r["max_t"], r["mask"] = ox, oy, dx, dy, max_t, mask
    return r


def world(api, bodies):
    w = sf.World(api, gravity_y=0.0)
    w.upload_bodies(bodies)
    return w


def test_ray_known_answers(ref_api):
    bodies = np.concatenate([
        body(SHAPE_CIRCLE, 5.0, 0.0, r=1.0),                       # 0
        body(SHAPE_RECT, 0.0, 5.0, hx=1.0, hy=0.5),                # 1 sharp box
        body(SHAPE_CAPSULE, -5.0, 0.0, hx=1.0, r=0.5),             # 2 along x
        body(SHAPE_RECT, 0.0, -5.0, hx=1.0, hy=1.0, r=0.5),        # 3 rounded box
        body(SHAPE_CIRCLE, 9.0, 0.0, r=1.0, category=4),           # 4 behind 0, other layer
    ])
    w = world(ref_api, bodies)
    rays = np.concatenate([
        ray(0, 0, 1, 0),                   # circle 0 at x = 4
        ray(0, 0, 0, 2),                   # box 1 bottom face at y = 4.5, direction length 2
        ray(0, 0, -1, 0),                  # capsule 2 end cap at x = -3.5
        ray(-5, 3, 0, -1),                 # capsule 2 side at y = 0.5
        ray(3, -8, -1, 1),                 # rounded corner of 3: centre (1,-6), r 0.5
        ray(5.2, 0.1, 1, 0),               # inside circle 0
        ray(0, 0, 1, 0, max_t=3.5),        # too short
        ray(0, 0, 1, 0, mask=4),           # only layer 4: passes 0, hits 4 at x = 8
        ray(0, 0, 1, 1, max_t=50),         # nothing on the diagonal
    ])
    h = w.ray_cast(rays)
    assert (h[0]["body"], h[0]["t"]) == (0, pytest.approx(4.0, abs=1e-6))
    assert (h[0]["nx"], h[0]["ny"]) == pytest.approx((-1.0, 0.0), abs=1e-6)
    assert (h[1]["body"], h[1]["t"]) == (1, pytest.approx(2.25, abs=1e-6))
    assert (h[1]["nx"], h[1]["ny"]) == pytest.approx((0.0, -1.0), abs=1e-6)
    assert (h[2]["body"], h[2]["t"]) == (2, pytest.approx(3.5, abs=1e-6))
    assert (h[3]["body"], h[3]["t"]) == (2, pytest.approx(2.5, abs=1e-6))
    assert (h[3]["nx"], h[3]["ny"]) == pytest.approx((0.0, 1.0), abs=1e-6)
    # ray from (3,-8) along (-1,1) passes through (1,-6)'s disc: enters at distance 0.5 before it
    t4 = 2.0 - 0.5 / math.sqrt(2.0)
    assert (h[4]["body"], h[4]["t"]) == (3, pytest.approx(t4, abs=1e-5))
    s = math.sqrt(0.5)
    assert (h[4]["nx"], h[4]["ny"]) == pytest.approx((s, -s), abs=1e-5)
    assert (h[5]["body"], h[5]["t"]) == (0, 0.0)
    assert (h[5]["nx"], h[5]["ny"]) == pytest.approx((-1.0, 0.0), abs=1e-6)
    assert h[6]["body"] == NONE
    assert (h[7]["body"], h[7]["t"]) == (4, pytest.approx(8.0, abs=1e-6))
    assert h[8]["body"] == NONE


def test_query_aabb_known_answers(ref_api):
    bodies = np.concatenate([
        body(SHAPE_CIRCLE, 0.0, 0.0, r=1.0),
        body(SHAPE_RECT, 3.0, 0.0, angle=math.pi / 4, hx=1.0, hy=1.0),  # rotated: half extent 1.414
        body(SHAPE_CAPSULE, 0.0, 4.0, hx=1.0, r=0.25, category=2),
    ])
    w = world(ref_api, bodies)
    assert w.query_aabb((0.9, -0.1), (1.7, 0.1)).tolist() == [0, 1]   # 3 - 1.414 = 1.586 < 1.7
    assert w.query_aabb((1.1, -0.1), (1.5, 0.1)).tolist() == []
    assert w.query_aabb((-5, -5), (5, 5)).tolist() == [0, 1, 2]
    assert w.query_aabb((-5, -5), (5, 5), mask=2).tolist() == [2]


def probe(shape, x, y, angle=0.0, hx=0.0, hy=0.0, r=0.0):
    q = np.zeros(1, dtype=_ffi.BODY_DESC_DTYPE)
    q["shape"], q["x"], q["y"], q["angle"], q["hx"], q["hy"], q["radius"] = shape, x, y, angle, hx, hy, r
    return q


def test_query_shape_and_point_known_answers(ref_api):
    bodies = np.concatenate([
        body(SHAPE_CIRCLE, 0.0, 0.0, r=1.0),                                    # 0
        body(SHAPE_RECT, 3.0, 0.0, angle=math.pi / 4, hx=1.0, hy=1.0),          # 1 diamond, tip at x = 1.586
        body(SHAPE_CAPSULE, 0.0, 4.0, hx=1.0, r=0.25, category=2),              # 2 from (-1,4) to (1,4)
        body(SHAPE_RECT, 10.0, 0.0, hx=1.0, hy=0.5, r=0.25),                    # 3 rounded box
    ])
    w = world(ref_api, bodies)
    # points
    assert w.query_point(0.5, 0.5).tolist() == [0]
    assert w.query_point(0.8, 0.8).tolist() == []              # 1.131 from the centre
    assert w.query_point(1.7, 0.0).tolist() == [1]             # inside the diamond's tip
    assert w.query_point(1.7, 0.2).tolist() == []              # beside the tip (|dy| < x - 1.586 fails)
    assert w.query_point(1.1, 4.15).tolist() == [2]            # in the cap: 0.18 from (1, 4)
    assert w.query_point(1.2, 4.2).tolist() == []              # outside it: 0.283 > 0.25
    assert w.query_point(1.1, 4.15, mask=1).tolist() == []     # the capsule is on layer 2
    assert w.query_point(11.15, 0.65).tolist() == [3]          # in the rounded corner: 0.212 from (11, .5)
    assert w.query_point(11.2, 0.7).tolist() == []             # outside it: 0.283
    # shapes
    assert w.query_shape(probe(SHAPE_CIRCLE, 1.3, 0.0, r=0.35)).tolist() == [0, 1]   # bridges the gap
    assert w.query_shape(probe(SHAPE_CIRCLE, 1.3, 0.0, r=0.25)).tolist() == []       # fits into it
    assert w.query_shape(probe(SHAPE_RECT, 0.0, 2.4, hx=1.5, hy=1.3)).tolist() == []     # y 1.1 .. 3.7, x up to 1.5
    assert w.query_shape(probe(SHAPE_RECT, 0.0, 2.4, hx=1.5, hy=1.45)).tolist() == [0, 2]
    assert w.query_shape(probe(SHAPE_CAPSULE, 5.0, 0.0, angle=math.pi / 2, hx=3.0, r=0.1)).tolist() == []
    assert w.query_shape(probe(SHAPE_CAPSULE, 6.5, 0.0, hx=2.4, r=0.2)).tolist() == [1, 3]
    # a bad probe is refused
    bad = probe(SHAPE_CIRCLE, 0.0, 0.0, r=-1.0)
    with pytest.raises(sf.SfpError):
        w.query_shape(bad)


def test_query_shape_against_independent_geometry(ref_api):
    """the overlap answer against separations computed WITHOUT the shared math header (float64
    support functions, tests/indep_geometry.py); probes whose separation is within 1e-4 of zero are
    left out"""
    import indep_geometry as ig
    rng = np.random.default_rng(21)
    b, _ = scenes.random_soup(160, 9, extent=6.0)
    b["angle"] = rng.uniform(-3.0, 3.0, len(b))
    w = world(ref_api, b)
    checked = hits = 0
    for k in range(60):
        shape = int(rng.integers(0, 3))
        hx = float(rng.uniform(0.1, 0.8)) if shape != SHAPE_CIRCLE else 0.0
        hy = float(rng.uniform(0.1, 0.5)) if shape == SHAPE_RECT else 0.0
        r = float(rng.uniform(0.0, 0.3)) if shape != SHAPE_CIRCLE else float(rng.uniform(0.0, 0.5))
        x, y, ang = float(rng.uniform(-6, 6)), float(rng.uniform(-6, 6)), float(rng.uniform(-3, 3))
        got = set(w.query_shape(probe(shape, x, y, ang, hx, hy, r)).tolist())
        VP = ig.to_world(ig.core_vertices(shape, hx, hy), x, y, ang)
        for i in range(len(b)):
            VB = ig.to_world(ig.core_vertices(int(b["shape"][i]), float(b["hx"][i]), float(b["hy"][i])),
                             float(b["x"][i]), float(b["y"][i]), float(b["angle"][i]))
            reach = np.hypot(b["hx"][i], b["hy"][i]) + b["radius"][i] + np.hypot(hx, hy) + r
            if np.hypot(b["x"][i] - x, b["y"][i] - y) > reach + 0.5:
                assert i not in got
                continue
            sep, _ = ig.separation(VB, float(b["radius"][i]), VP, r)
            if abs(sep) < 1e-4:
                continue
            checked += 1
            hits += sep < 0
            assert (i in got) == (sep < 0), (k, i, sep)
    assert checked > 300 and hits > 40


def random_rays(n, seed, extent_x, extent_y):
    rng = np.random.default_rng(seed)
    r = np.zeros(n, dtype=RAY_DTYPE)
    r["ox"] = rng.uniform(-extent_x, extent_x, n)
    r["oy"] = rng.uniform(-1.0, extent_y, n)
    ang = rng.uniform(0, 2 * np.pi, n)
    ln = rng.uniform(0.2, 3.0, n)
    r["dx"], r["dy"] = np.cos(ang) * ln, np.sin(ang) * ln
    r["max_t"] = rng.uniform(0.5, 40.0, n)
    r["mask"] = np.where(rng.uniform(size=n) < 0.8, 0xFFFF, 1 << rng.integers(0, 4, n))
    r[: n // 20]["dx"] = 0.0  # axis-aligned rays too
    r[n // 20: n // 10]["dy"] = 0.0
    return r


@pytest.mark.gpu
@pytest.mark.parametrize("scene", ["pile", "soup"])
def test_ray_cast_and_box_query_match_oracle(gpu_api, ref_api, scene):
    if scene == "pile":
        b, j = scenes.mixed_pile(4000)
        ex, ey = 26.0, 30.0
    else:
        b, j = scenes.random_soup(900, 5)
        ex, ey = 12.0, 22.0
    g, r = sf.World(gpu_api), sf.World(ref_api)
    for w in (g, r):
        w.upload_bodies(b)
        w.set_joints(j)
    rays = random_rays(3000, 11, ex, ey)
    for frames in (0, 40, 80):
        for _ in range(frames):
            g.step()
            r.step()
        hg, hr = g.ray_cast(rays), r.ray_cast(rays)
        assert (hr["body"] != NONE).sum() > 500
        assert hg.tobytes() == hr.tobytes()
        for lo, hi in (((-3.0, 0.0), (2.0, 4.0)), ((-50.0, -5.0), (50.0, 0.4)), ((5.0, 5.0), (5.1, 5.1))):
            assert np.array_equal(g.query_aabb(lo, hi), r.query_aabb(lo, hi))
        rng = np.random.default_rng(frames)
        found = 0
        for k in range(40):
            shape = int(rng.integers(0, 3))
            q = probe(shape, rng.uniform(-ex / 2, ex / 2), rng.uniform(0.0, 6.0), rng.uniform(-3, 3),
                      rng.uniform(0.1, 1.0) if shape else 0.0, rng.uniform(0.1, 0.6) if shape == 2 else 0.0,
                      rng.uniform(0.0, 0.4))
            a, c = g.query_shape(q, 0xFFFF if k % 4 else 1), r.query_shape(q, 0xFFFF if k % 4 else 1)
            assert np.array_equal(a, c)
            found += len(c)
            assert np.array_equal(g.query_point(float(q["x"][0]), float(q["y"][0])),
                                  r.query_point(float(q["x"][0]), float(q["y"][0])))
        assert found > 20
    # a query must not disturb the simulation: keep stepping, still bit-identical
    for _ in range(10):
        g.step()
        r.step()
    assert g.download_bodies().tobytes() == r.download_bodies().tobytes()
    assert np.array_equal(g.get_pairs(), r.get_pairs())


@pytest.mark.gpu
def test_long_rays_from_far_away_match_oracle(gpu_api, ref_api):
    """rays that start kilometres from the scene (tens of thousands of 0.5 m cells away, more
    with small cells) and rays that miss it: the device walk is confined to the bodies' cell
    bounding box, so its length does not depend on how long the ray is"""
    b, j = scenes.mixed_pile(3000)
    rng = np.random.default_rng(3)
    n = 400
    rays = np.zeros(n, dtype=RAY_DTYPE)
    ang = rng.uniform(0, 2 * np.pi, n)
    dist = rng.choice([30.0, 900.0, 9000.0, 60000.0], n)
    tx, ty = rng.uniform(-10, 10, n), rng.uniform(0.0, 6.0, n)   # aim at the pile (most rays)
    rays["ox"], rays["oy"] = tx + dist * np.cos(ang), ty + dist * np.sin(ang)
    ln = rng.uniform(0.5, 2.0, n)
    rays["dx"], rays["dy"] = -np.cos(ang) * ln, -np.sin(ang) * ln
    rays[:40]["dy"] += 0.7                                        # some miss the scene entirely
    rays["max_t"] = 1.0e6
    rays["mask"] = 0xFFFF
    for cell in (0.5, 0.05):
        g, r = sf.World(gpu_api, cell_size=cell), sf.World(ref_api, cell_size=cell)
        for w in (g, r):
            w.upload_bodies(b)
        for _ in range(30):
            g.step()
            r.step()
        hg, hr = g.ray_cast(rays), r.ray_cast(rays)
        assert (hr["body"] != NONE).sum() > 200
        assert hg.tobytes() == hr.tobytes()
