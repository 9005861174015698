"""General convex polygons (north_star: "rounded-polygon colliders"): 3..8 vertices from a side
table + a rounding radius.  CPU tier: closed-form checks on the oracle and input validation.
GPU tier: bit-exact parity of a pile that is half polygons, rays included."""
import math

import numpy as np
import pytest

import starframe_b200 as sf
from starframe_b200 import _ffi, scenes
from starframe_b200._ffi import (POLYGON_DTYPE, RAY_DTYPE, SHAPE_CIRCLE, SHAPE_POLYGON, SHAPE_RECT)

FIELDS = ("x", "y", "c", "s", "vx", "vy", "omega")


def poly_body(index, x, y, angle=0.0, r=0.0, **kw):
    b = scenes.empty_bodies(1)
    b["shape"] = SHAPE_POLYGON | (index << 8)
    b["x"], b["y"], b["angle"], b["radius"] = x, y, angle, r
    for k, v in kw.items():
        b[k] = v
    return b


def world(api, bodies, table, **params):
    w = sf.World(api, **params)
    w.set_polygons(table)
    w.upload_bodies(bodies)
    return w


def test_hexagon_rests_on_its_flat_side(ref_api):
    """regular hexagon R = 0.5 with a flat side down: centre settles at the apothem R cos 30"""
    table = scenes.regular_polygon(6, 0.5)  # vertices at 0, 60, ...: flat side down (240-300)
    floor = scenes.static_rect(0.0, -0.5, 5.0, 0.5)
    w = world(ref_api, np.concatenate([floor, poly_body(0, 0.0, 0.45)]), table)
    for _ in range(180):
        w.step()
    s = w.download_bodies()[1]
    assert s["y"] == pytest.approx(0.5 * math.cos(math.pi / 6), abs=2e-3)
    assert abs(s["angle"]) < 2e-3 and abs(s["vy"]) < 0.02
    c = w.get_contacts()
    assert len(c) == 1 and c[0]["count"] == 2  # face contact: two points
    assert sorted(c[0]["px"].tolist()) == pytest.approx([-0.25, 0.25], abs=2e-3)


def test_triangle_circle_face_contact(ref_api):
    """circle resting against the right face of an equilateral triangle: normal = face normal"""
    table = scenes.regular_polygon(3, 1.0)  # vertices at 0, 120, 240 degrees
    tri = poly_body(0, 0.0, 0.0, flags=_ffi.BODY_STATIC)
    # face between vertex 0 (1,0) and vertex 1 (-0.5, 0.866): outward normal at 60 degrees, at
    # distance 0.5 from the centre
    n = (math.cos(math.pi / 3), math.sin(math.pi / 3))
    ball = scenes.empty_bodies(1)
    ball["shape"], ball["radius"] = SHAPE_CIRCLE, 0.2
    ball["x"], ball["y"] = n[0] * 0.68, n[1] * 0.68  # 0.5 + 0.2 - 0.02
    w = world(ref_api, np.concatenate([tri, ball]), table, gravity_y=0.0)
    w.step(1e-4, 1)
    c = w.get_contacts()
    assert len(c) == 1 and c[0]["count"] == 1
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx(n, abs=1e-5)
    assert c[0]["depth"][0] == pytest.approx(0.02, abs=1e-5)


def test_rounded_polygon_vertex_contact_and_mass(ref_api):
    """two rounded squares given as polygons touch vertex to vertex; and a polygon square has
    the mass of the equivalent rectangle (falls, collides and splits corrections equally)"""
    sq = scenes.polygon([(-0.4, -0.4), (0.4, -0.4), (0.4, 0.4), (-0.4, 0.4)])
    a = poly_body(0, 0.0, 0.0, r=0.1, flags=_ffi.BODY_STATIC)
    b = poly_body(0, 0.93, 0.93, r=0.1)
    w = world(ref_api, np.concatenate([a, b]), sq, gravity_y=0.0)
    w.step(1e-4, 1)
    c = w.get_contacts()
    s = math.sqrt(0.5)
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((s, s), abs=1e-5)
    assert c[0]["depth"][0] == pytest.approx(0.2 - 0.13 * math.sqrt(2), abs=1e-6)
    # same scene with the built-in rounded rect gives the same manifold
    ra = scenes.empty_bodies(2)
    ra["shape"], ra["hx"], ra["hy"], ra["radius"] = SHAPE_RECT, 0.4, 0.4, 0.1
    ra["x"], ra["y"] = [0.0, 0.93], [0.0, 0.93]
    ra["flags"][0] = _ffi.BODY_STATIC
    w2 = sf.World(ref_api, gravity_y=0.0)
    w2.upload_bodies(ra)
    w2.step(1e-4, 1)
    c2 = w2.get_contacts()
    assert c2[0]["depth"][0] == pytest.approx(c[0]["depth"][0], abs=1e-6)


def test_ray_hits_polygon_faces_and_rounded_corners(ref_api):
    table = scenes.regular_polygon(4, math.sqrt(2.0), phase=math.pi / 4)  # square [-1,1]^2
    w = world(ref_api, np.concatenate([poly_body(0, 5.0, 0.0), poly_body(0, 0.0, 5.0, r=0.5)]), table,
              gravity_y=0.0)
    rays = np.zeros(3, dtype=RAY_DTYPE)
    rays["max_t"], rays["mask"] = 100.0, 0xFFFF
    rays["dx"] = [1.0, 0.0, -1.0]
    rays["dy"] = [0.0, 1.0, 1.0]
    rays["ox"] = [0.0, 0.0, 3.0]
    rays["oy"] = [0.0, 0.0, 2.0]
    h = w.ray_cast(rays)
    assert (h[0]["body"], h[0]["t"]) == (0, pytest.approx(4.0, abs=1e-5))
    assert (h[1]["body"], h[1]["t"]) == (1, pytest.approx(3.5, abs=1e-5))  # 5 - 1 - 0.5
    # diagonal ray towards the rounded corner at (1, 4): disc of radius 0.5 around it
    t = 2.0 - 0.5 / math.sqrt(2.0)
    assert (h[2]["body"], h[2]["t"]) == (1, pytest.approx(t, abs=1e-5))


def test_invalid_polygons_and_indices_are_rejected(ref_api):
    w = sf.World(ref_api)
    cw = scenes.regular_polygon(5, 0.3)
    cw["x"][0, :5] = cw["x"][0, :5][::-1].copy()
    cw["y"][0, :5] = cw["y"][0, :5][::-1].copy()
    concave = np.zeros(1, dtype=POLYGON_DTYPE)
    concave["count"] = 5
    concave["x"][0, :5] = [-1, 1, 1, 0, -1]
    concave["y"][0, :5] = [-1, -1, 1, 0.2, 1]
    off = scenes.regular_polygon(4, 0.3)
    off["x"] += 0.2
    two = scenes.regular_polygon(3, 0.3)
    two["count"] = 2
    for bad in (cw, concave, off, two):
        with pytest.raises(sf.SfpError) as e:
            w.set_polygons(bad)
        assert e.value.code == _ffi.SFP_E_INVALID
    w.set_polygons(scenes.polygon_table())
    with pytest.raises(sf.SfpError):
        w.upload_bodies(poly_body(99, 0.0, 1.0))  # index beyond the table
    w.upload_bodies(poly_body(3, 0.0, 1.0))


def test_polygon_pile_settles(ref_api):
    b, j, table = scenes.polygon_pile(500)
    w = world(ref_api, b, table)
    e_prev = None
    for f in range(420):
        w.step()
        if f >= 300 and f % 30 == 0:
            st = w.stats()
            e = st["kinetic_energy"] + st["potential_energy"]
            assert e_prev is None or e <= e_prev + 2e-3 * abs(e_prev)
            e_prev = e
    st = w.stats()
    assert st["max_penetration"] < 0.02 and st["kinetic_energy"] < 0.5
    s = w.download_bodies()
    assert s["y"][3:].min() > 0.0 and np.isfinite(s["x"]).all()


@pytest.mark.gpu
def test_polygon_pile_gpu_matches_oracle(gpu_api, ref_api):
    b, j, table = scenes.polygon_pile(1500)
    g, r = world(gpu_api, b, table), world(ref_api, b, table)
    rng = np.random.default_rng(2)
    rays = np.zeros(1500, dtype=RAY_DTYPE)
    rays["ox"], rays["oy"] = rng.uniform(-16, 16, 1500), rng.uniform(-1, 30, 1500)
    ang = rng.uniform(0, 2 * np.pi, 1500)
    rays["dx"], rays["dy"], rays["max_t"], rays["mask"] = np.cos(ang), np.sin(ang), 30.0, 0xFFFF
    for block in range(5):
        for _ in range(30):
            g.step()
            r.step()
        assert np.array_equal(g.get_pairs(), r.get_pairs())
        sg, sr = g.stats(), r.stats()
        for k in ("n_pairs", "n_contacts", "n_manifolds", "n_colours", "colour_rounds"):
            assert sg[k] == sr[k], (k, sg[k], sr[k])
        a, c = g.download_bodies(), r.download_bodies()
        for f in FIELDS:
            assert np.array_equal(a[f].view(np.uint32), c[f].view(np.uint32)), (block, f)
        assert g.ray_cast(rays).tobytes() == r.ray_cast(rays).tobytes()
        # box, point and shape queries: the polygons answer by their vertices; probe 3 + k is
        # itself a general polygon (an index into the world's table)
        found = 0
        for k in range(24):
            x, y = float(rng.uniform(-8, 8)), float(rng.uniform(0.2, 8.0))
            assert np.array_equal(g.query_aabb((x, y), (x + 1.0, y + 0.6)), r.query_aabb((x, y), (x + 1.0, y + 0.6)))
            assert np.array_equal(g.query_point(x, y), r.query_point(x, y))
            q = np.zeros(1, dtype=_ffi.BODY_DESC_DTYPE)
            q["x"], q["y"], q["angle"], q["radius"] = x, y, float(rng.uniform(-3, 3)), 0.05
            q["shape"] = SHAPE_POLYGON | ((k % len(table)) << 8)
            a, c = g.query_shape(q), r.query_shape(q)
            assert np.array_equal(a, c)
            found += len(c)
        assert found > 10
    assert g.get_contacts().tobytes() == r.get_contacts().tobytes()
    assert sg["n_contacts"] > 1500


@pytest.mark.gpu
def test_polygon_world_multi_kernel_mode(gpu_api, ref_api):
    b, j, table = scenes.polygon_pile(600)
    g = world(gpu_api, b, table, flags=_ffi.FLAG_MULTI_KERNEL)
    r = world(ref_api, b, table)
    for _ in range(25):
        g.step()
        r.step()
    assert g.download_bodies().tobytes() == r.download_bodies().tobytes()


@pytest.mark.gpu
def test_batch_rejects_polygons_without_a_table(gpu_api):
    b = sf.Batch(gpu_api, 1, 4)
    bodies = np.concatenate([poly_body(0, float(i), 1.0) for i in range(4)])
    with pytest.raises(sf.SfpError):
        b.upload(bodies)


@pytest.mark.gpu
def test_batched_polygon_worlds_match_oracle(gpu_api, ref_api):
    """general polygons inside batched worlds (sfp_batch_set_polygons + the polygon kernel variant)"""
    nw, n, frames = 4, 300, 90
    worlds, table = [], None
    for w in range(nw):
        b, j, table = scenes.polygon_pile(n, seed=scenes.BASE_SEED + 60 + w, cols=20)
        worlds.append(b)
    bodies = np.stack(worlds)
    batch = sf.Batch(gpu_api, nw, bodies.shape[1])
    batch.set_polygons(table)
    batch.upload(bodies)
    refs = []
    for w in range(nw):
        r = sf.World(ref_api)
        r.set_polygons(table)
        r.upload_bodies(bodies[w])
        refs.append(r)
    for f in range(frames):
        batch.step()
        for r in refs:
            r.step()
    out, st = batch.download(), batch.stats()
    for w in range(nw):
        ro = refs[w].download_bodies()
        for f in ("x", "y", "c", "s", "vx", "vy", "omega"):
            assert np.array_equal(out[w][f].view(np.uint32), ro[f].view(np.uint32)), (w, f)
        rs = refs[w].stats()
        for k in ("n_pairs", "n_contacts", "n_manifolds", "n_colours"):
            assert st[k][w] == rs[k], (w, k, st[k][w], rs[k])
    assert st["n_contacts"].sum() > nw * n // 2
