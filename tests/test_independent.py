"""Checks that do NOT go through include/sfp_math.h ("the oracle and the kernels
share their arithmetic"): the narrowphase and the XPBD position solve of the library against
float64 numpy restatements that share no code and no algorithm with it (tests/indep_geometry.py:
signed distance functions on dense boundary samplings instead of SAT + clipping; Mueller et al.
2020's contact update written out).

CPU tier: the oracle.  GPU tier: the CUDA library through the C ABI, same cases.
Tolerances (stated, float32 library against float64 sampling): depth 1e-3 m, normal 1 degree,
positions after one projection 2e-5 m, angles 6e-5 rad + 0.5 % of the turn.
"""
import numpy as np
import pytest

import starframe_b200 as sf
from starframe_b200 import scenes
from starframe_b200._ffi import SHAPE_POLYGON

import indep_geometry as ig

KINDS = (ig.CIRCLE, ig.CAPSULE, ig.RECT, ig.POLYGON)
TYPE_PAIRS = [(a, b) for i, a in enumerate(KINDS) for b in KINDS[i:]]  # all 10
PER_TYPE_PAIR = 120
PITCH = 4.0


def random_shape(rng, kind, table):
    d = dict(kind=kind, hx=0.0, hy=0.0, r=0.0, poly=None, pidx=0)
    if kind == ig.CIRCLE:
        d["r"] = rng.uniform(0.05, 0.25)
    elif kind == ig.CAPSULE:
        d["hx"], d["r"] = rng.uniform(0.05, 0.3), rng.uniform(0.05, 0.12)
    elif kind == ig.RECT:
        d["hx"], d["hy"] = rng.uniform(0.06, 0.3), rng.uniform(0.06, 0.3)
        d["r"] = rng.choice([0.0, rng.uniform(0.01, 0.06)])
    else:
        d["pidx"] = int(rng.integers(0, 6))  # (the sliver, index 6, is thinner than the depths used)
        d["poly"] = table[d["pidx"]]
        d["r"] = rng.choice([0.0, rng.uniform(0.01, 0.05)])
    for k in ("hx", "hy", "r"):
        d[k] = float(np.float32(d[k]))
    return d


def cases(seed, lo, hi):
    """pairs of shapes placed so that the independent separation is a target in [lo, hi]"""
    rng = np.random.default_rng(seed)
    table = scenes.polygon_table()
    out = []
    for ta, tb in TYPE_PAIRS:
        for _ in range(PER_TYPE_PAIR):
            A, B = random_shape(rng, ta, table), random_shape(rng, tb, table)
            A["angle"], B["angle"] = (float(np.float32(rng.uniform(-np.pi, np.pi))) for _ in range(2))
            th = rng.uniform(0, 2 * np.pi)
            u = np.array([np.cos(th), np.sin(th)])
            target = rng.uniform(lo, hi)
            VA = ig.to_world(ig.core_vertices(A["kind"], A["hx"], A["hy"], A["poly"]), 0.0, 0.0, A["angle"])
            vb = ig.core_vertices(B["kind"], B["hx"], B["hy"], B["poly"])
            t0, t1 = 0.0, 2.5
            for _ in range(40):
                t = 0.5 * (t0 + t1)
                s, _n = ig.separation(VA, A["r"], ig.to_world(vb, *(t * u), B["angle"]), B["r"],
                                      directions=720)
                if s < target:
                    t0 = t
                else:
                    t1 = t
            B["off"] = (0.5 * (t0 + t1)) * u
            out.append((A, B))
    return out


def build_world(api, pairs):
    table = scenes.polygon_table()
    b = scenes.empty_bodies(2 * len(pairs))
    cols = 40
    for i, (A, B) in enumerate(pairs):
        ox, oy = PITCH * (i % cols), PITCH * (i // cols)
        for j, S in enumerate((A, B)):
            q = b[2 * i + j]
            off = S.get("off", np.zeros(2))
            q["x"], q["y"], q["angle"] = ox + off[0], oy + off[1], S["angle"]
            q["hx"], q["hy"], q["radius"] = S["hx"], S["hy"], S["r"]
            q["shape"] = S["kind"] if S["kind"] != ig.POLYGON else (SHAPE_POLYGON | (S["pidx"] << 8))
            q["density"] = 1.0 + 0.5 * j
    w = sf.World(api, gravity_x=0.0, gravity_y=0.0)
    w.set_polygons(table)
    w.upload_bodies(b)
    return w, b


def independent(b, i, pairs):
    """separation and normal of pair i from the float32 inputs the library was given"""
    out = []
    for j, S in enumerate(pairs[i]):
        q = b[2 * i + j]
        V = ig.to_world(ig.core_vertices(S["kind"], float(q["hx"]), float(q["hy"]), S["poly"]),
                        float(q["x"]), float(q["y"]), float(q["angle"]))
        out.append((V, float(q["radius"])))
    return ig.separation(out[0][0], out[0][1], out[1][0], out[1][1])


def check_narrowphase(api):
    touching = cases(20261107, -0.015, 0.006)
    w, b = build_world(api, touching)
    w.step(1.0 / 480.0, 1)
    w.sync()
    con = {(int(c["a"]), int(c["b"])): c for c in w.get_contacts()}
    worst_d, worst_ang, checked_n = 0.0, 0.0, 0
    for i in range(len(touching)):
        sep, n = independent(b, i, touching)
        assert (2 * i, 2 * i + 1) in con, (i, sep, touching[i])
        c = con[(2 * i, 2 * i + 1)]
        k = int(c["count"])
        depth = float(np.max(c["depth"][:k]))
        worst_d = max(worst_d, abs(depth + sep))
        assert abs(depth + sep) <= 1e-3, (i, depth, -sep, touching[i])
        # the normal of a rounded corner against a rounded corner turns quickly with the pose when
        # the two core features are close together; compare where it is well conditioned
        VA = ig.to_world(ig.core_vertices(touching[i][0]["kind"], float(b[2 * i]["hx"]), float(b[2 * i]["hy"]),
                                          touching[i][0]["poly"]), float(b[2 * i]["x"]), float(b[2 * i]["y"]),
                         float(b[2 * i]["angle"]))
        core_gap = sep + float(b[2 * i]["radius"]) + float(b[2 * i + 1]["radius"])
        if core_gap > 0.02 or k == 2:
            ang = np.degrees(np.arccos(np.clip(n[0] * c["nx"] + n[1] * c["ny"], -1.0, 1.0)))
            worst_ang = max(worst_ang, ang)
            checked_n += 1
            assert ang <= 1.0, (i, ang, n, (c["nx"], c["ny"]), touching[i])
        del VA
    assert checked_n > len(touching) // 2
    apart = cases(20261108, 0.014, 0.05)
    w2, b2 = build_world(api, apart)
    w2.step(1.0 / 480.0, 1)
    w2.sync()
    assert len(w2.get_contacts()) == 0
    return worst_d, worst_ang


def test_narrowphase_against_float64_sampling_oracle(ref_api):
    worst_d, worst_ang = check_narrowphase(ref_api)
    print(f"worst depth error {worst_d:.2e} m, worst normal error {worst_ang:.3f} deg")


@pytest.mark.gpu
def test_narrowphase_against_float64_sampling_gpu(gpu_api):
    check_narrowphase(gpu_api)


# ---- XPBD position projection -----------------------------------------------------------------
def check_projection(api):
    rng = np.random.default_rng(7)
    table = scenes.polygon_table()
    shapes = (ig.CIRCLE, ig.CAPSULE, ig.RECT)
    pairs = []
    for _ in range(300):
        A, B = (random_shape(rng, shapes[int(rng.integers(0, 3))], table) for _ in range(2))
        A["angle"], B["angle"] = (float(np.float32(rng.uniform(-np.pi, np.pi))) for _ in range(2))
        pairs.append((A, B))
    # place with the sampling separation at -2 .. -8 mm
    rng2 = np.random.default_rng(8)
    # face-to-face rectangles (two-point manifolds with both points penetrating), placed directly
    aligned = []
    for _ in range(60):
        A, B = random_shape(rng, ig.RECT, table), random_shape(rng, ig.RECT, table)
        th = float(np.float32(rng.uniform(-np.pi, np.pi)))
        A["angle"] = th
        B["angle"] = float(np.float32(th + rng.uniform(-0.004, 0.004)))
        nrm, tan = np.array([np.cos(th), np.sin(th)]), np.array([-np.sin(th), np.cos(th)])
        gap = A["hx"] + A["r"] + B["hx"] + B["r"] - rng.uniform(0.003, 0.007)
        B["off"] = gap * nrm + rng.uniform(-0.3, 0.3) * min(A["hy"], B["hy"]) * tan
        aligned.append((A, B))
    for A, B in pairs:
        th = rng2.uniform(0, 2 * np.pi)
        u = np.array([np.cos(th), np.sin(th)])
        target = rng2.uniform(-0.008, -0.002)
        VA = ig.to_world(ig.core_vertices(A["kind"], A["hx"], A["hy"]), 0.0, 0.0, A["angle"])
        vb = ig.core_vertices(B["kind"], B["hx"], B["hy"])
        t0, t1 = 0.0, 2.5
        for _ in range(40):
            t = 0.5 * (t0 + t1)
            s, _n = ig.separation(VA, A["r"], ig.to_world(vb, *(t * u), B["angle"]), B["r"], directions=720)
            t0, t1 = (t, t1) if s < target else (t0, t)
        B["off"] = (0.5 * (t0 + t1)) * u
    pairs = pairs + aligned
    w, b = build_world(api, pairs)
    before = w.download_bodies()
    w.step(1.0 / 480.0, 1)
    w.sync()
    after = w.download_bodies()
    con = {(int(c["a"]), int(c["b"])): c for c in w.get_contacts()}
    n1 = n2 = 0
    for i in range(len(pairs)):
        c = con.get((2 * i, 2 * i + 1))
        assert c is not None
        k = int(c["count"])
        n = np.array([float(c["nx"]), float(c["ny"])])
        sa, sb = before[2 * i], before[2 * i + 1]
        xa, xb = np.array([sa["x"], sa["y"]], dtype=np.float64), np.array([sb["x"], sb["y"]], dtype=np.float64)
        tha, thb = float(np.arctan2(sa["s"], sa["c"])), float(np.arctan2(sb["s"], sb["c"]))
        ima, iia = ig.inverse_mass(pairs[i][0]["kind"], *(float(b[2 * i][f]) for f in ("hx", "hy", "radius", "density")))
        imb, iib = ig.inverse_mass(pairs[i][1]["kind"], *(float(b[2 * i + 1][f]) for f in ("hx", "hy", "radius", "density")))
        # depth and normal as reported for narrowphase time (checked against independent geometry
        # above); the lever arms from the reported contact point, which is given at the END of
        # the substep and moves with the bodies: carried back with each body's own motion
        mid = [np.array([float(c["px"][q]), float(c["py"][q])]) for q in range(k)]
        dep = [float(c["depth"][q]) for q in range(k)]

        def rot_of(s0, s1):
            da = np.arctan2(s1["s"], s1["c"]) - np.arctan2(s0["s"], s0["c"])
            return np.array([[np.cos(da), -np.sin(da)], [np.sin(da), np.cos(da)]])

        # pa0 on A's surface and pb0 = pa0 - depth * n on B's at the start; each moves rigidly
        # with its body (motions T_A, T_B known from the two downloads) and the library reports
        # mid = (T_A pa0 + T_B pb0) / 2: linear in pa0, solved exactly
        RA, RB = rot_of(sa, after[2 * i]), rot_of(sb, after[2 * i + 1])
        xa1 = np.array([after[2 * i]["x"], after[2 * i]["y"]], dtype=np.float64)
        xb1 = np.array([after[2 * i + 1]["x"], after[2 * i + 1]["y"]], dtype=np.float64)
        ras, rbs = [], []
        for q in range(k):
            rhs = 2.0 * mid[q] - xa1 - xb1 + RA @ xa + RB @ (dep[q] * n + xb)
            pa0 = np.linalg.solve(RA + RB, rhs)
            ras.append(pa0 - xa)
            rbs.append(pa0 - dep[q] * n - xb)
        if k == 1:
            if dep[0] <= 0.0:
                continue
            res = ig.xpbd_contact_1pt(xa, tha, ima, iia, xb, thb, imb, iib, ras[0], rbs[0], n, dep[0])
            n1 += 1
        else:
            if min(dep) <= 0.0:
                continue  # one point inside the margin only: a different active set
            res = ig.xpbd_contact_2pt(xa, tha, ima, iia, xb, thb, imb, iib, ras, rbs, n, dep)
            if res is None:
                continue
            n2 += 1
        exp_xa, exp_tha, exp_xb, exp_thb = res
        for got, ex, eth, th0 in ((after[2 * i], exp_xa, exp_tha, tha), (after[2 * i + 1], exp_xb, exp_thb, thb)):
            assert abs(got["x"] - ex[0]) <= 2e-5 and abs(got["y"] - ex[1]) <= 2e-5, (i, k, got, ex)
            dth = (np.arctan2(got["s"], got["c"]) - eth + np.pi) % (2 * np.pi) - np.pi
            # the library turns a body by normalise(q + dtheta perp q), i.e. by atan(dtheta): the
            # linearised update of the paper, third order away from the angle itself
            # (0.5 % of the turn: the lever arms come from a float32 contact midpoint)
            assert abs(dth) <= 6e-5 + 5e-3 * abs(eth - th0) + 0.5 * abs(eth - th0) ** 3, (i, k, dth, eth - th0)
    assert n1 >= 100 and n2 >= 15, (n1, n2)


def test_position_projection_against_float64_xpbd_oracle(ref_api):
    check_projection(ref_api)


@pytest.mark.gpu
def test_position_projection_against_float64_xpbd_gpu(gpu_api):
    check_projection(gpu_api)
