"""Property tests with generated worlds (hypothesis, derandomised so every run sees the same
examples): small ADVERSARIAL scenes — bodies placed anywhere, deep overlaps, coincident centres,
slivers next to discs, every shape kind, static and dynamic, several layers, random joints —
rather than the orderly piles of the other files.

CPU tier: invariants of the oracle that hold for any correct implementation of the path.
GPU tier: the CUDA path against the oracle on the same worlds, bit for bit.
"""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

import starframe_b200 as sf
from starframe_b200 import scenes
from starframe_b200._ffi import (BODY_STATIC, JOINT_DESC_DTYPE, SHAPE_CAPSULE, SHAPE_CIRCLE,
                                 SHAPE_POLYGON, SHAPE_RECT)

FIELDS = ("x", "y", "c", "s", "vx", "vy", "omega")
SETTINGS = dict(derandomize=True, deadline=None,
                suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])

def f32(lo, hi):
    """floats of width 32 between the float32 neighbours of lo and hi"""
    return st.floats(float(np.float32(lo)), float(np.float32(hi)), width=32, allow_nan=False,
                     allow_infinity=False)

body = st.fixed_dictionaries({
    "kind": st.sampled_from([SHAPE_CIRCLE, SHAPE_CAPSULE, SHAPE_RECT, SHAPE_POLYGON]),
    # positions on a coarse lattice with a small offset: coincident and barely-touching bodies are
    # likely, which is where degenerate normals and contact / no-contact decisions live
    "gx": st.integers(-4, 4), "gy": st.integers(0, 5),
    "ox": f32(-0.05, 0.05), "oy": f32(-0.05, 0.05),
    "angle": f32(-3.2, 3.2),
    "hx": f32(0.02, 0.45), "hy": f32(0.02, 0.45), "r": f32(0.0, 0.2),
    "vx": f32(-3.0, 3.0), "vy": f32(-3.0, 3.0), "omega": f32(-6.0, 6.0),
    "static": st.booleans(), "layer": st.integers(0, 2), "poly": st.integers(0, 6),
    "mu": f32(0.0, 1.0), "e": f32(0.0, 0.8),
})


def build(bodies_spec, joints_spec):
    n = len(bodies_spec)
    b = scenes.empty_bodies(n)
    for i, s in enumerate(bodies_spec):
        k = s["kind"]
        b[i]["x"] = 0.5 * s["gx"] + s["ox"]
        b[i]["y"] = 0.5 * s["gy"] + s["oy"]
        b[i]["angle"] = s["angle"]
        b[i]["vx"], b[i]["vy"], b[i]["omega"] = s["vx"], s["vy"], s["omega"]
        b[i]["mu_s"], b[i]["mu_d"], b[i]["restitution"] = s["mu"], 0.7 * s["mu"], s["e"]
        b[i]["flags"] = BODY_STATIC if s["static"] else 0
        b[i]["category"] = 1 << s["layer"]
        b[i]["mask"] = 0xFFFF if s["layer"] < 2 else 0x3
        if k == SHAPE_CIRCLE:
            b[i]["shape"], b[i]["radius"] = k, max(s["r"], 0.03)
        elif k == SHAPE_CAPSULE:
            b[i]["shape"], b[i]["hx"], b[i]["radius"] = k, s["hx"], max(s["r"], 0.02)
        elif k == SHAPE_RECT:
            b[i]["shape"], b[i]["hx"], b[i]["hy"], b[i]["radius"] = k, s["hx"], s["hy"], s["r"]
        else:
            b[i]["shape"], b[i]["radius"] = SHAPE_POLYGON | (s["poly"] << 8), min(s["r"], 0.05)
    floor = scenes.static_rect(0.0, -0.6, 4.0, 0.5)
    b = np.concatenate([b, floor])
    j = np.zeros(len(joints_spec), dtype=JOINT_DESC_DTYPE)
    keep = []
    for q, (a, c, rest, kind, comp) in enumerate(joints_spec):
        a, c = a % n, c % n
        if a == c:
            continue
        j[q]["body_a"], j[q]["body_b"] = a, c
        j[q]["rest"], j[q]["kind"], j[q]["compliance"] = rest, kind, comp
        keep.append(q)
    return b, j[keep]


worlds = st.tuples(
    st.lists(body, min_size=1, max_size=24),
    st.lists(st.tuples(st.integers(0, 63), st.integers(0, 63), f32(0.2, 1.5), st.integers(0, 1),
                       st.sampled_from([0.0, 1.0e-4])), max_size=6),
    st.sampled_from([0.25, 0.5, 1.5]),   # cell size
    st.sampled_from([1, 4, 8]),          # substeps
)


def make(api, b, j, cell):
    w = sf.World(api, cell_size=cell)
    w.set_polygons(scenes.polygon_table())
    w.upload_bodies(b)
    w.set_joints(j)
    return w


@settings(max_examples=200, **SETTINGS)
@given(worlds)
def test_oracle_invariants_on_generated_worlds(ref_api, spec):
    bodies_spec, joints_spec, cell, substeps = spec
    b, j = build(bodies_spec, joints_spec)
    w = make(ref_api, b, j, cell)
    before = w.download_bodies()
    brute = w.brute_pairs(1 / 60)
    w.step(1 / 60, substeps)
    # broadphase: exactly the brute-force set, sorted, a < b, unique
    got = w.get_pairs()
    assert np.array_equal(got, brute)
    assert np.all(got["a"] < got["b"])
    # colouring: proper over dynamic bodies (the serial group excepted)
    _, pc, jc = w.get_colours()
    dyn = (b["flags"] & BODY_STATIC) == 0
    used = {}
    items = [(int(p["a"]), int(p["b"]), int(c)) for p, c in zip(got, pc)]
    items += [(int(q["body_a"]), int(q["body_b"]), int(c)) for q, c in zip(j, jc)]
    for a, c_, col in items:
        if col == 63:
            continue
        for body_ in (a, c_):
            if dyn[body_]:
                assert (body_, col) not in used, (body_, col)
                used[(body_, col)] = True
    # state: finite, rotors unit, statics untouched
    after = w.download_bodies()
    for f in FIELDS:
        assert np.all(np.isfinite(after[f])), f
    assert np.allclose(after["c"] ** 2 + after["s"] ** 2, 1.0, atol=1e-5)
    for f in FIELDS:
        assert np.array_equal(after[f][~dyn], before[f][~dyn]), f


@pytest.mark.gpu
@settings(max_examples=300, **SETTINGS)
@given(worlds)
def test_gpu_matches_oracle_on_generated_worlds(gpu_api, ref_api, spec):
    bodies_spec, joints_spec, cell, substeps = spec
    b, j = build(bodies_spec, joints_spec)
    g, r = make(gpu_api, b, j, cell), make(ref_api, b, j, cell)
    for _ in range(3):
        g.step(1 / 60, substeps)
        r.step(1 / 60, substeps)
    assert np.array_equal(g.get_pairs(), r.get_pairs())
    sg, sr = g.stats(), r.stats()
    for k in ("n_pairs", "n_contacts", "n_manifolds", "n_colours", "colour_rounds"):
        assert sg[k] == sr[k], (k, sg[k], sr[k])
    a, c = g.download_bodies(), r.download_bodies()
    for f in FIELDS + ("angle",):
        assert np.array_equal(a[f].view(np.uint32), c[f].view(np.uint32)), f


batch_worlds = st.tuples(
    st.lists(st.tuples(st.lists(body, min_size=12, max_size=12),
                       st.lists(st.tuples(st.integers(0, 63), st.integers(0, 63), f32(0.2, 1.5),
                                          st.integers(0, 1), st.sampled_from([0.0, 1.0e-4])),
                                max_size=4)),
             min_size=2, max_size=4),
    st.sampled_from([0.25, 0.5]),
)


@pytest.mark.gpu
@settings(max_examples=100, **SETTINGS)
@given(batch_worlds)
def test_gpu_batch_matches_oracle_on_generated_worlds(gpu_api, ref_api, spec):
    """the same adversarial worlds, several of them in one batch (polygons and joints: the full
    kernel variant), each against its own oracle world"""
    specs, cell = spec
    built = [build(bs, js) for bs, js in specs]
    nw, n = len(built), len(built[0][0])
    jmax = max(len(j) for _, j in built)
    bodies = np.stack([b for b, _ in built])
    batch = sf.Batch(gpu_api, nw, n, cell_size=cell)
    batch.set_polygons(scenes.polygon_table())
    batch.upload(bodies)
    if jmax:
        packed = np.zeros((nw, jmax), dtype=JOINT_DESC_DTYPE)
        counts = np.zeros(nw, dtype=np.uint32)
        for w, (_, j) in enumerate(built):
            packed[w, :len(j)] = j
            counts[w] = len(j)
        batch.set_joints(packed, counts)
    refs = [make(ref_api, b, j, cell) for b, j in built]
    for _ in range(3):
        batch.step()
        for r in refs:
            r.step()
    out, st_ = batch.download(), batch.stats()
    for w, r in enumerate(refs):
        c = r.download_bodies()
        for f in FIELDS:
            assert np.array_equal(out[w][f].view(np.uint32), c[f].view(np.uint32)), (w, f)
        rs = r.stats()
        for k in ("n_pairs", "n_contacts", "n_colours"):
            assert st_[k][w] == rs[k], (w, k)


ray = st.tuples(st.integers(-10, 10), st.integers(-3, 12), f32(-0.01, 0.01), f32(-0.01, 0.01),
                st.sampled_from([-2.0, -1.0, -0.5, 0.0, 0.5, 1.0, 2.0]),
                st.sampled_from([-2.0, -1.0, -0.5, 0.0, 0.5, 1.0, 2.0]),
                st.sampled_from([0.5, 3.0, 50.0]), st.sampled_from([0xFFFF, 0x1, 0x6]))


@pytest.mark.gpu
@settings(max_examples=150, **SETTINGS)
@given(worlds, st.lists(ray, min_size=1, max_size=24))
def test_gpu_queries_match_oracle_on_generated_worlds(gpu_api, ref_api, spec, rays_spec):
    """ray casts (cell walk on the device, brute force in the oracle) and box queries on the
    adversarial worlds; ray origins sit on or next to cell boundaries, directions are often
    axis-aligned — the cases where a grid walk goes wrong"""
    from starframe_b200._ffi import RAY_DTYPE

    bodies_spec, joints_spec, cell, substeps = spec
    b, j = build(bodies_spec, joints_spec)
    g, r = make(gpu_api, b, j, cell), make(ref_api, b, j, cell)
    g.step(1 / 60, substeps)
    r.step(1 / 60, substeps)
    rays = np.zeros(len(rays_spec), dtype=RAY_DTYPE)
    for k, (ix, iy, ox, oy, dx, dy, tmax, mask) in enumerate(rays_spec):
        rays[k]["ox"], rays[k]["oy"] = 0.25 * ix + ox, 0.25 * iy + oy
        rays[k]["dx"], rays[k]["dy"] = dx, dy
        rays[k]["max_t"], rays[k]["mask"] = tmax, mask
    hg, hr = g.ray_cast(rays), r.ray_cast(rays)
    assert hg.tobytes() == hr.tobytes(), (hg, hr)
    for lo, hi in (((-1.0, 0.0), (1.0, 1.0)), ((-5.0, -1.0), (5.0, 4.0)), ((0.3, 0.3), (0.3, 0.3))):
        assert np.array_equal(g.query_aabb(lo, hi), r.query_aabb(lo, hi))
