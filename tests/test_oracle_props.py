"""Structural invariants of the CPU oracle: they hold for ANY correct
implementation of the path, so they check the oracle independently of the CUDA code."""
import numpy as np
import pytest

import starframe_b200 as sf
from starframe_b200 import _ffi, scenes
from starframe_b200._ffi import BODY_STATIC

FIELDS = ("x", "y", "c", "s", "vx", "vy", "omega")


def world(api, bodies, joints=None, **params):
    w = sf.World(api, **params)
    w.upload_bodies(bodies)
    w.set_joints(scenes.no_joints() if joints is None else joints)
    return w


@pytest.mark.parametrize("seed", range(6))
@pytest.mark.parametrize("cell", [0.25, 0.5, 1.7])
def test_hash_broadphase_equals_brute_force(ref_api, seed, cell):
    """the spatial hash must report exactly the O(n^2) AABB-overlap + filter set: sorted,
    a < b, no duplicates — for any cell size (bodies much larger and much smaller than a cell)"""
    b, _ = scenes.random_soup(400, 100 + seed, extent=4.0)
    w = world(ref_api, b, cell_size=cell)
    brute = w.brute_pairs(1 / 60)
    w.step(1 / 60, 1)
    got = w.get_pairs()
    assert np.all(got["a"] < got["b"])
    key = got["a"].astype(np.uint64) << np.uint64(32) | got["b"].astype(np.uint64)
    assert np.all(np.diff(key.astype(np.int64)) > 0)  # strictly increasing: sorted + unique
    assert np.array_equal(got, brute)


def test_broadphase_far_from_origin_and_clamped_cells(ref_api):
    """cells are clamped to a 65536^2 lattice; bodies beyond it must still pair correctly"""
    b, _ = scenes.random_soup(200, 7, extent=3.0)
    b["x"] += 20000.0  # cell index 40000 > 32767 at cell 0.5 -> all clamp into the border column
    w = world(ref_api, b)
    brute = w.brute_pairs(1 / 60)
    w.step(1 / 60, 1)
    assert np.array_equal(w.get_pairs(), brute)


def test_static_static_pairs_are_dropped(ref_api):
    b, _ = scenes.random_soup(100, 3, extent=1.0)
    b["flags"] = BODY_STATIC
    w = world(ref_api, b)
    w.step()
    assert len(w.get_pairs()) == 0


@pytest.mark.parametrize("scene", ["soup", "pile", "ropes"])
def test_colouring_is_proper(ref_api, scene):
    """no two constraints of one colour share a dynamic body; colours are 0..n_colours-1"""
    if scene == "soup":
        b, j = scenes.random_soup(500, 5, pitch=0.5)
    elif scene == "pile":
        b, j = scenes.mixed_pile(1500)
    else:
        b, j = scenes.ropes(n_ropes=30, links=24, n_free=200)
    w = world(ref_api, b, j)
    for _ in range(40 if scene == "pile" else 3):
        w.step()
    pairs, pc, jc = w.get_colours()
    ncol = w.stats()["n_colours"]
    dyn = (b["flags"] & BODY_STATIC) == 0
    used = set()
    for (a, bb), c in zip(pairs, pc):
        assert c < ncol
        for body in (a, bb):
            if dyn[body]:
                assert (body, c) not in used
                used.add((body, c))
    for jd, c in zip(j, jc):
        assert c < ncol
        for body in (jd["body_a"], jd["body_b"]):
            if dyn[body]:
                assert (body, c) not in used
                used.add((body, c))
    assert set(np.unique(np.concatenate([pc, jc])).tolist()) == set(range(ncol))


def test_static_bodies_never_move(ref_api):
    b, j = scenes.random_soup(400, 21)
    w = world(ref_api, b, j)
    s0 = w.download_bodies()
    for _ in range(30):
        w.step()
    s1 = w.download_bodies()
    st = (b["flags"] & BODY_STATIC) != 0
    assert st.sum() > 10
    for f in FIELDS:
        assert np.array_equal(s0[f][st], s1[f][st])


def test_momentum_conserved_without_statics_and_gravity(ref_api):
    """every projection applies equal and opposite impulses: total linear and angular momentum
    of an isolated system stay put"""
    b, _ = scenes.random_soup(300, 31, pitch=1.0)  # no initial overlap: gentle collisions
    b = b[:-1]  # drop the floor
    b["flags"] = 0
    b["mask"] = 0xFFFF
    b["category"] = 1
    w = world(ref_api, b, gravity_y=0.0)

    def momentum():
        s = w.download_bodies()
        m = np.empty(len(b))
        inertia = np.empty(len(b))
        for i, d in enumerate(b):
            r, hx, hy = float(d["radius"]), float(d["hx"]), float(d["hy"])
            if d["shape"] == 0:
                m[i] = np.pi * r * r
                inertia[i] = 0.5 * m[i] * r * r
            elif d["shape"] == 1:
                mb, mc = 4 * hx * r, np.pi * r * r
                m[i] = mb + mc
                inertia[i] = mb * (4 * hx * hx + 4 * r * r) / 12 + mc * (
                    0.5 * r * r + hx * hx + 2 * hx * 4 * r / (3 * np.pi))
            else:
                ww, hh = 2 * (hx + r), 2 * (hy + r)
                m[i] = ww * hh
                inertia[i] = m[i] * (ww * ww + hh * hh) / 12
        px = (m * s["vx"]).sum()
        py = (m * s["vy"]).sum()
        L = (inertia * s["omega"] + m * (s["x"] * s["vy"] - s["y"] * s["vx"])).sum()
        scale = (m * np.hypot(s["vx"], s["vy"])).sum()
        return np.array([px, py, L]), scale

    p0, scale = momentum()
    touched = 0
    for _ in range(60):
        w.step()
        touched += w.stats()["n_contacts"]
    p1, _ = momentum()
    assert touched > 100
    assert np.abs(p1[:2] - p0[:2]).max() < 2e-4 * scale
    # angular momentum about the origin: exact for the impulses, first-order in the positional
    # corrections (arms move while a penetration is resolved)
    assert abs(p1[2] - p0[2]) < 1e-2 * scale


def test_colour_order_vs_reference_order_single_substep(ref_api):
    """north_star: single-substep positions / velocities within a stated tolerance that accounts
    for Gauss-Seidel ORDER (reference order vs colouring).  The GPU matches the colour-order
    oracle bit for bit (tests/test_gpu_parity.py), so the whole tolerance is the ordering term,
    measured here oracle-vs-oracle on a settled 1200-body pile:
      - discrete results (pairs, contact count) identical;
      - positions: a body whose contacts are solved in a different order differs by up to the
        size of one contact correction (penetrations in a resting pile are a few mm), so the
        bound is absolute: max < 10 mm, median < 0.5 mm (measured: max 5.9 mm, median 0.14 mm;
        for scale, gravity alone moves a body g h^2 = 0.043 mm per substep);
      - velocities are displacement / h with h = 1/480 s, so the same term reads 480x larger."""
    b, j = scenes.mixed_pile(1200)
    wc = world(ref_api, b, j, solver_order=_ffi.ORDER_COLOUR)
    for _ in range(300):
        wc.step()
    state = wc.download_bodies()
    wr = world(ref_api, b, j, solver_order=_ffi.ORDER_REFERENCE)
    wc2 = world(ref_api, b, j, solver_order=_ffi.ORDER_COLOUR)
    a = wc2.step_host(state, 1 / 480, 1)
    r = wr.step_host(state, 1 / 480, 1)
    assert np.array_equal(wc2.get_pairs(), wr.get_pairs())
    assert wc2.stats()["n_contacts"] == wr.stats()["n_contacts"]
    dpos = np.hypot(a["x"] - r["x"], a["y"] - r["y"])
    assert dpos.max() < 1e-2
    assert np.median(dpos) < 5e-4
    dvel = np.hypot(a["vx"] - r["vx"], a["vy"] - r["vy"])
    assert dvel.max() < 1e-2 * 480
    assert np.median(dvel) < 5e-4 * 480


def test_oracle_is_deterministic(ref_api):
    b, j = scenes.mixed_pile(800)
    outs = []
    for _ in range(2):
        w = world(ref_api, b, j)
        for _ in range(40):
            w.step()
        outs.append(w.download_bodies().tobytes())
    assert outs[0] == outs[1]


def test_narrowphase_is_symmetric_under_swap(ref_api):
    """swapping the two bodies of a pair flips the normal and keeps depth / point"""
    b, _ = scenes.random_soup(2, 1)
    rng = np.random.default_rng(5)
    checked = 0
    for _ in range(300):
        pair = scenes.empty_bodies(2)
        pair["shape"] = rng.integers(0, 3, 2)
        pair["hx"] = np.where(pair["shape"] == 0, 0.0, rng.uniform(0.1, 0.4, 2))
        pair["hy"] = np.where(pair["shape"] == 2, rng.uniform(0.1, 0.3, 2), 0.0)
        pair["radius"] = rng.uniform(0.05, 0.15, 2)
        pair["x"] = [0.0, rng.uniform(-0.6, 0.6)]
        pair["y"] = [0.0, rng.uniform(-0.6, 0.6)]
        pair["angle"] = rng.uniform(-3, 3, 2)
        res = []
        for order in ([0, 1], [1, 0]):
            w = world(ref_api, pair[order], gravity_y=0.0)
            w.step(1e-5, 1)
            res.append(w.get_contacts())
        c0, c1 = res
        assert len(c0) == len(c1)
        if len(c0) == 0 or c0[0]["count"] != c1[0]["count"]:
            continue
        if c0[0]["count"] == 1:
            assert c0[0]["nx"] == pytest.approx(-c1[0]["nx"], abs=2e-3)
            assert c0[0]["ny"] == pytest.approx(-c1[0]["ny"], abs=2e-3)
            assert c0[0]["depth"][0] == pytest.approx(c1[0]["depth"][0], abs=2e-4)
            checked += 1
    assert checked > 30


def test_invalid_inputs_rejected(ref_api):
    w = sf.World(ref_api)
    b = scenes.empty_bodies(2)
    b["radius"] = 0.1
    b["density"][1] = 0.0  # dynamic body without mass
    with pytest.raises(sf.SfpError) as e:
        w.upload_bodies(b)
    assert e.value.code == _ffi.SFP_E_INVALID
    b["density"][1] = 1.0
    b["x"][0] = np.nan
    with pytest.raises(sf.SfpError):
        w.upload_bodies(b)
    b["x"][0] = 0.0
    w.upload_bodies(b)
    j = np.zeros(1, dtype=_ffi.JOINT_DESC_DTYPE)
    j["body_a"], j["body_b"] = 0, 5  # out of range
    with pytest.raises(sf.SfpError):
        w.set_joints(j)
    with pytest.raises(sf.SfpError):
        w.step(0.0, 8)
    with pytest.raises(sf.SfpError):
        w.step(1 / 60, 0)


def test_rollout_metrics_pile(ref_api):
    """long rollout judged on penetration, energy and rest (north_star): a 600-body pile
    settles — penetration stays small, total energy never grows, the pile comes to rest"""
    b, j = scenes.mixed_pile(600)
    w = world(ref_api, b, j)
    e_prev, worst_pen = None, 0.0
    for f in range(480):
        w.step()
        if f % 30 == 29:
            st = w.stats()
            e = st["kinetic_energy"] + st["potential_energy"]
            if e_prev is not None:
                assert e <= e_prev + 1e-3 * abs(e_prev) + 1e-3
            e_prev = e
            if f > 240:
                worst_pen = max(worst_pen, st["max_penetration"])
    st = w.stats()
    assert worst_pen < 0.01
    assert st["kinetic_energy"] < 1e-3 * abs(st["potential_energy"]) + 0.05


def test_serial_group_when_a_body_has_more_partners_than_colours(ref_api):
    """a dynamic plank carrying 100 bodies: 63 proper colours do not suffice; the remaining
    constraints join the serial group (colour 63) and the step still works and stays put"""
    b, j = scenes.crowded_plank(100)
    w = world(ref_api, b, j)
    for _ in range(120):
        w.step()
    pairs, pc, _ = w.get_colours()
    st = w.stats()
    assert st["n_colours"] == 64 and st["overflow"] == 0
    plank = 1
    on_plank = (pairs["a"] == plank) | (pairs["b"] == plank)
    assert on_plank.sum() > 63
    serial = pc == 63
    assert serial.sum() >= on_plank.sum() - 63
    # proper colours are proper: the plank appears at most once per colour below 63
    cols = pc[on_plank & ~serial]
    assert np.unique(cols).size == cols.size
    s = w.download_bodies()
    assert abs(s["y"][plank] - 0.06) < 5e-3 and abs(s["angle"][plank]) < 1e-3
    assert np.all(s["y"][2:] > 0.1) and np.abs(s["vy"]).max() < 0.05  # riders rest on the plank


def test_threaded_colour_order_step_has_the_same_bits(ref_api, monkeypatch):
    """SFP_REF_THREADS: constraints of one colour touch disjoint dynamic bodies and every other
    loop is per body or per pair, so the thread count must not change a single bit"""
    import starframe_b200 as sf
    from starframe_b200 import scenes

    b, j = scenes.ropes(n_ropes=12, links=16, n_free=900)
    out = {}
    for threads in (1, 3):
        monkeypatch.setenv("SFP_REF_THREADS", str(threads))
        w = sf.World(ref_api)
        w.upload_bodies(b)
        w.set_joints(j)
        for _ in range(25):
            w.step()
        out[threads] = (w.download_bodies(), w.stats())
    for f in ("x", "y", "c", "s", "vx", "vy", "omega", "angle"):
        assert np.array_equal(out[1][0][f].view(np.uint32), out[3][0][f].view(np.uint32)), f
    for k in ("n_pairs", "n_contacts", "n_manifolds", "n_colours"):
        assert out[1][1][k] == out[3][1][k]
    assert out[1][1]["n_contacts"] > 300
