"""Known-answer tests that pin the CPU oracle to PHYSICS (closed-form answers), because it cannot
be pinned to the reference: the reference tree, its tests and fixtures are not mounted
(SURVEY.md §8c).  Every expected value below is derived by hand in the test.
"""
import math

import numpy as np
import pytest

import starframe_b200 as sf
from starframe_b200 import _ffi
from starframe_b200._ffi import (BODY_STATIC, JOINT_DESC_DTYPE, JOINT_DISTANCE, JOINT_ROPE,
                                 SHAPE_CAPSULE, SHAPE_CIRCLE, SHAPE_RECT)
from starframe_b200.scenes import empty_bodies, no_joints, static_rect

G = 9.81


def world(ref_api, bodies, joints=None, **params):
    w = sf.World(ref_api, **params)
    w.upload_bodies(bodies)
    w.set_joints(no_joints() if joints is None else joints)
    return w


def body(shape, x, y, angle=0.0, hx=0.0, hy=0.0, r=0.0, static=False, **kw):
    b = empty_bodies(1)
    b["shape"], b["x"], b["y"], b["angle"] = shape, x, y, angle
    b["hx"], b["hy"], b["radius"] = hx, hy, r
    b["flags"] = BODY_STATIC if static else 0
    for k, v in kw.items():
        b[k] = v
    return b


def cat(*bs):
    return np.concatenate(bs)


# ---------------------------------------------------------------------------------------------
def test_free_fall_semi_implicit_euler(ref_api):
    """No contacts: v_n = v0 + n h g, y_n = y0 + h sum_k v_k = y0 + n h v0 + g h^2 n(n+1)/2."""
    b = body(SHAPE_CIRCLE, 0.0, 10.0, r=0.5, vy=1.0, vx=2.0)
    w = world(ref_api, b)
    frames, S, dt = 30, 8, 1 / 60
    for _ in range(frames):
        w.step(dt, S)
    n, h = frames * S, dt / S
    s = w.download_bodies()[0]
    assert s["vy"] == pytest.approx(1.0 - n * h * G, rel=1e-5)
    assert s["y"] == pytest.approx(10.0 + n * h * 1.0 - G * h * h * n * (n + 1) / 2, rel=1e-5)
    assert s["x"] == pytest.approx(2.0 * n * h, rel=1e-5)
    assert s["vx"] == pytest.approx(2.0, rel=1e-6)


def test_free_spin_keeps_angular_velocity(ref_api):
    """integrate -> derive must return omega exactly (rot_advance / rot_delta_tan are inverses);
    the angle advances by atan(h w) per substep."""
    b = body(SHAPE_RECT, 0.0, 0.0, hx=0.3, hy=0.2, omega=5.0)
    w = world(ref_api, b, gravity_y=0.0)
    w.step(1 / 60, 8)
    s = w.download_bodies()[0]
    assert s["omega"] == pytest.approx(5.0, rel=2e-6)
    assert s["angle"] == pytest.approx(8 * math.atan(5.0 / 480), rel=1e-5)
    assert s["c"] ** 2 + s["s"] ** 2 == pytest.approx(1.0, abs=1e-6)


def one_contact(ref_api, a, b, **params):
    """contact manifold of two bodies after one tiny zero-gravity substep"""
    w = world(ref_api, cat(a, b), gravity_y=0.0, **params)
    w.step(1e-4, 1)
    return w, w.get_contacts()


def test_circle_circle_depth_normal_and_equal_split(ref_api):
    a = body(SHAPE_CIRCLE, 0.0, 0.0, r=0.5)
    b = body(SHAPE_CIRCLE, 0.9, 0.0, r=0.5)
    w, c = one_contact(ref_api, a, b)
    assert len(c) == 1 and c[0]["count"] == 1
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((1.0, 0.0), abs=1e-6)
    assert c[0]["depth"][0] == pytest.approx(0.1, abs=1e-6)
    s = w.download_bodies()
    # equal masses: each moves half the depth, apart
    assert s["x"][0] == pytest.approx(-0.05, abs=1e-6)
    assert s["x"][1] == pytest.approx(0.95, abs=1e-6)


def test_mass_ratio_split(ref_api):
    """density 4:1 with equal size -> inverse masses 1:4 -> displacement 1/5 : 4/5 of the depth"""
    a = body(SHAPE_CIRCLE, 0.0, 0.0, r=0.5, density=4.0)
    b = body(SHAPE_CIRCLE, 0.9, 0.0, r=0.5, density=1.0)
    w, _ = one_contact(ref_api, a, b)
    s = w.download_bodies()
    assert s["x"][0] == pytest.approx(-0.02, abs=1e-6)
    assert s["x"][1] == pytest.approx(0.98, abs=1e-6)


def test_circle_vs_static_only_dynamic_moves(ref_api):
    a = body(SHAPE_CIRCLE, 0.0, 0.0, r=0.5, static=True)
    b = body(SHAPE_CIRCLE, 0.0, 0.8, r=0.5)
    w, c = one_contact(ref_api, a, b)
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((0.0, 1.0), abs=1e-6)
    s = w.download_bodies()
    assert (s["x"][0], s["y"][0]) == (0.0, 0.0)
    assert s["y"][1] == pytest.approx(1.0, abs=1e-6)


def test_circle_capsule_side_and_cap(ref_api):
    cap = body(SHAPE_CAPSULE, 0.0, 0.0, hx=1.0, r=0.2, static=True)
    # above the middle of the segment: normal straight up, depth = 0.2+0.3-0.45
    _, c = one_contact(ref_api, cap, body(SHAPE_CIRCLE, 0.3, 0.45, r=0.3))
    assert c[0]["count"] == 1
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((0.0, 1.0), abs=1e-6)
    assert c[0]["depth"][0] == pytest.approx(0.05, abs=1e-6)
    # beyond the end cap: normal along the line from the cap centre (1,0) to the circle centre
    _, c = one_contact(ref_api, cap, body(SHAPE_CIRCLE, 1.3, 0.3, r=0.3))
    d = math.hypot(0.3, 0.3)
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((0.3 / d, 0.3 / d), abs=1e-6)
    assert c[0]["depth"][0] == pytest.approx(0.5 - d, abs=1e-6)


def test_circle_rect_face_and_corner(ref_api):
    box = body(SHAPE_RECT, 0.0, 0.0, hx=1.0, hy=0.5, r=0.0, static=True)
    _, c = one_contact(ref_api, box, body(SHAPE_CIRCLE, 0.2, 0.75, r=0.3))
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((0.0, 1.0), abs=1e-6)
    assert c[0]["depth"][0] == pytest.approx(0.05, abs=1e-6)
    # corner (1, 0.5): centre at (1.2, 0.7) -> distance 0.2828, depth 0.3-0.2828
    _, c = one_contact(ref_api, box, body(SHAPE_CIRCLE, 1.2, 0.7, r=0.3))
    d = math.hypot(0.2, 0.2)
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((0.2 / d, 0.2 / d), abs=1e-5)
    assert c[0]["depth"][0] == pytest.approx(0.3 - d, abs=1e-6)


def test_circle_inside_rect_core_uses_face_normal(ref_api):
    box = body(SHAPE_RECT, 0.0, 0.0, hx=1.0, hy=0.5, static=True)
    _, c = one_contact(ref_api, box, body(SHAPE_CIRCLE, 0.0, 0.45, r=0.1))
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((0.0, 1.0), abs=1e-6)
    assert c[0]["depth"][0] == pytest.approx(0.15, abs=1e-6)


def test_rect_on_rect_two_points(ref_api):
    """box (1 x 1) resting 0.01 deep in a wide static box: two contact points at the box's
    bottom corners, normal up, depth 0.01 each"""
    ground = body(SHAPE_RECT, 0.0, -0.5, hx=5.0, hy=0.5, static=True)
    box = body(SHAPE_RECT, 0.3, 0.49, hx=0.5, hy=0.5)
    _, c = one_contact(ref_api, ground, box)
    assert len(c) == 1 and c[0]["count"] == 2
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((0.0, 1.0), abs=1e-6)
    assert sorted(c[0]["px"].tolist()) == pytest.approx([-0.2, 0.8], abs=1e-5)
    assert c[0]["depth"].tolist() == pytest.approx([0.01, 0.01], abs=1e-5)


def test_rounded_rect_corner_contact(ref_api):
    """two rounded boxes meeting corner to corner along the diagonal: vertex-vertex normal"""
    a = body(SHAPE_RECT, 0.0, 0.0, hx=0.4, hy=0.4, r=0.1, static=True)
    b = body(SHAPE_RECT, 0.93, 0.93, hx=0.4, hy=0.4, r=0.1)
    _, c = one_contact(ref_api, a, b)
    assert c[0]["count"] == 1
    s = math.sqrt(0.5)
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((s, s), abs=1e-5)
    # core corners (0.4,0.4) and (0.53,0.53): distance 0.13*sqrt2, minus the two radii
    assert c[0]["depth"][0] == pytest.approx(0.2 - 0.13 * math.sqrt(2), abs=1e-6)


def test_capsule_capsule_parallel_two_points(ref_api):
    a = body(SHAPE_CAPSULE, 0.0, 0.0, hx=1.0, r=0.2, static=True)
    b = body(SHAPE_CAPSULE, 0.5, 0.35, hx=1.0, r=0.2)
    _, c = one_contact(ref_api, a, b)
    assert c[0]["count"] == 2
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((0.0, 1.0), abs=1e-6)
    assert c[0]["depth"].tolist() == pytest.approx([0.05, 0.05], abs=1e-6)
    assert sorted(c[0]["px"].tolist()) == pytest.approx([-0.5, 1.0], abs=1e-5)


def test_capsule_capsule_crossed_single_point(ref_api):
    a = body(SHAPE_CAPSULE, 0.0, 0.0, hx=1.0, r=0.2, static=True)
    b = body(SHAPE_CAPSULE, 0.0, 1.3, angle=math.pi / 2, hx=1.0, r=0.2)
    _, c = one_contact(ref_api, a, b)
    # b's lower end (0, 0.3) above a's axis: separation 0.3 - 0.4
    assert c[0]["count"] == 1
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((0.0, 1.0), abs=1e-5)
    assert c[0]["depth"][0] == pytest.approx(0.1, abs=1e-5)


def test_capsule_on_rect_two_points(ref_api):
    ground = body(SHAPE_RECT, 0.0, -0.5, hx=5.0, hy=0.5, static=True)
    cap = body(SHAPE_CAPSULE, 0.0, 0.19, hx=0.5, r=0.2)
    _, c = one_contact(ref_api, ground, cap)
    assert c[0]["count"] == 2
    assert (c[0]["nx"], c[0]["ny"]) == pytest.approx((0.0, 1.0), abs=1e-6)
    assert sorted(c[0]["px"].tolist()) == pytest.approx([-0.5, 0.5], abs=1e-5)
    assert c[0]["depth"].tolist() == pytest.approx([0.01, 0.01], abs=1e-5)


def test_separated_beyond_margin_no_contact(ref_api):
    a = body(SHAPE_CIRCLE, 0.0, 0.0, r=0.5)
    b = body(SHAPE_CIRCLE, 1.0105, 0.0, r=0.5)  # gap 0.0105 > contact_margin 0.01
    _, c = one_contact(ref_api, a, b)
    assert len(c) == 0
    b2 = body(SHAPE_CIRCLE, 1.005, 0.0, r=0.5)  # gap 0.005 < margin: speculative point, inactive
    w, c = one_contact(ref_api, a, b2)
    assert len(c) == 1 and c[0]["depth"][0] == pytest.approx(-0.005, abs=1e-6)
    assert c[0]["lambda_n"][0] == 0.0
    s = w.download_bodies()
    assert s["x"][0] == 0.0 and s["x"][1] == pytest.approx(1.005)


def test_layers_filter(ref_api):
    a = body(SHAPE_CIRCLE, 0.0, 0.0, r=0.5, category=1, mask=2)
    b = body(SHAPE_CIRCLE, 0.5, 0.0, r=0.5, category=2, mask=2)  # b does not accept a's layer 1
    _, c = one_contact(ref_api, a, b)
    assert len(c) == 0
    b["mask"] = 1
    _, c = one_contact(ref_api, a, b)
    assert len(c) == 1


# ---------------------------------------------------------------------------------------------
def joint(a, b, rest, kind=JOINT_DISTANCE, compliance=0.0, anchor_a=(0, 0), anchor_b=(0, 0)):
    j = np.zeros(1, dtype=JOINT_DESC_DTYPE)
    j["body_a"], j["body_b"], j["rest"], j["kind"], j["compliance"] = a, b, rest, kind, compliance
    j["ax"], j["ay"] = anchor_a
    j["bx"], j["by"] = anchor_b
    return j


def test_distance_joint_split_by_inverse_mass(ref_api):
    """|x2-x1| = 1.3, rest 1.0, inverse masses 1:4 -> corrections 0.06 and 0.24 towards each other"""
    a = body(SHAPE_CIRCLE, 0.0, 0.0, r=0.1, density=4.0, category=1, mask=0)
    b = body(SHAPE_CIRCLE, 1.3, 0.0, r=0.1, density=1.0, category=1, mask=0)
    w = world(ref_api, cat(a, b), joint(0, 1, 1.0), gravity_y=0.0)
    w.step(1 / 480, 1)
    s = w.download_bodies()
    assert s["x"][0] == pytest.approx(0.06, abs=1e-6)
    assert s["x"][1] == pytest.approx(1.06, abs=1e-6)
    # and the velocity derived from the correction: dx / h
    assert s["vx"][0] == pytest.approx(0.06 * 480, rel=1e-4)


def test_rope_joint_is_unilateral(ref_api):
    a = body(SHAPE_CIRCLE, 0.0, 0.0, r=0.1, category=1, mask=0, static=True)
    b = body(SHAPE_CIRCLE, 0.7, 0.0, r=0.1, category=1, mask=0)
    w = world(ref_api, cat(a, b), joint(0, 1, 1.0, JOINT_ROPE), gravity_y=0.0)
    w.step(1 / 480, 1)
    assert w.download_bodies()["x"][1] == pytest.approx(0.7)  # slack: untouched
    b["x"] = 1.2
    w = world(ref_api, cat(a, b), joint(0, 1, 1.0, JOINT_ROPE), gravity_y=0.0)
    w.step(1 / 480, 1)
    assert w.download_bodies()["x"][1] == pytest.approx(1.0, abs=1e-6)  # taut: pulled back


def test_compliant_joint(ref_api):
    """alpha~ = alpha/h^2; d_lambda = -c/(w1+w2+alpha~); b is the only mover (a static)."""
    a = body(SHAPE_CIRCLE, 0.0, 0.0, r=0.1, category=1, mask=0, static=True)
    b = body(SHAPE_CIRCLE, 1.5, 0.0, r=0.1, category=1, mask=0, density=1.0)
    alpha, h = 1e-5, 1 / 480
    w = world(ref_api, cat(a, b), joint(0, 1, 1.0, compliance=alpha), gravity_y=0.0)
    w.step(h, 1)
    wb = 1.0 / (math.pi * 0.01)
    expect = 1.5 - 0.5 * wb / (wb + alpha / h ** 2)
    assert w.download_bodies()["x"][1] == pytest.approx(expect, rel=1e-5)


def test_pendulum_keeps_length_and_energy(ref_api):
    a = body(SHAPE_CIRCLE, 0.0, 0.0, r=0.05, category=1, mask=0, static=True)
    b = body(SHAPE_CIRCLE, 1.0, 0.0, r=0.05, category=1, mask=0)
    w = world(ref_api, cat(a, b), joint(0, 1, 1.0))
    e0 = None
    for f in range(120):
        w.step()
        s = w.download_bodies()[1]
        st = w.stats()
        e = st["kinetic_energy"] + st["potential_energy"]
        e0 = e if e0 is None else e0
        assert math.hypot(s["x"], s["y"]) == pytest.approx(1.0, abs=2e-3)
    m = math.pi * 0.05 ** 2
    # the position projection dissipates (first-order method; measured loss 8.4 % of m g L over
    # 2 s at 8 substeps); it must never GAIN energy
    assert e <= e0 + 1e-4 * m * G
    assert e0 - e < 0.15 * m * G


# ---------------------------------------------------------------------------------------------
def test_ball_rests_on_floor(ref_api):
    floor = static_rect(0.0, -0.5, 5.0, 0.5)
    ball = body(SHAPE_CIRCLE, 0.0, 0.52, r=0.5)
    w = world(ref_api, cat(floor, ball))
    for _ in range(120):
        w.step()
    s = w.download_bodies()[1]
    assert s["y"] == pytest.approx(0.5, abs=2e-3)
    assert abs(s["vy"]) < 0.05
    assert w.stats()["max_penetration"] < 2e-3


@pytest.mark.parametrize("e", [0.0, 0.5, 0.8])
def test_restitution_bounce_speed(ref_api, e):
    """impact speed ~ sqrt(2 g drop) is far above the 2|g|h threshold -> rebound = e * impact"""
    floor = static_rect(0.0, -0.5, 5.0, 0.5, e=e)
    ball = body(SHAPE_CIRCLE, 0.0, 2.5, r=0.5, restitution=e)
    w = world(ref_api, cat(floor, ball))
    prev_vy, impact, rebound = 0.0, None, None
    for _ in range(200):
        w.step()
        vy = float(w.download_bodies()["vy"][1])
        if impact is None and vy > prev_vy + 1.0:  # velocity jumped upward: the bounce
            impact, rebound = prev_vy, vy
            break
        prev_vy = vy
    assert impact is not None
    assert impact == pytest.approx(-math.sqrt(2 * G * 2.0), rel=0.05)
    # one frame of gravity (g*dt) acts between the two samples
    assert rebound == pytest.approx(-e * impact, abs=0.06 * abs(impact) + G / 60)


@pytest.mark.parametrize("mu_s,holds", [(0.6, True), (0.2, False)])
def test_static_friction_on_incline(ref_api, mu_s, holds):
    """tan(20 deg) = 0.364: a box holds iff mu_s >= tan(theta)"""
    th = math.radians(20.0)
    ramp = static_rect(0.0, 0.0, 10.0, 0.5, angle=th, mu_s=mu_s, mu_d=0.9 * mu_s)
    n = (-math.sin(th), math.cos(th))
    d = 0.5 + 0.25 + 0.001
    box = body(SHAPE_RECT, n[0] * d, n[1] * d, angle=th, hx=0.25, hy=0.25, mu_s=mu_s,
               mu_d=0.9 * mu_s)
    w = world(ref_api, cat(ramp, box))
    for _ in range(30):
        w.step()
    s0 = w.download_bodies()[1].copy()
    for _ in range(120):
        w.step()
    s1 = w.download_bodies()[1]
    slid = math.hypot(s1["x"] - s0["x"], s1["y"] - s0["y"])
    if holds:
        assert slid < 1e-3  # no creep at all inside the cone
    else:
        # slides with a = g (sin - mu_d cos) over 2 s (+ the half second before s0)
        a = G * (math.sin(th) - 0.9 * mu_s * math.cos(th))
        expect = 0.5 * a * (2.5 ** 2 - 0.5 ** 2)
        assert slid == pytest.approx(expect, rel=0.25)


def test_dynamic_friction_deceleration(ref_api):
    """box sliding on a flat floor decelerates at mu_d * g"""
    mu_d = 0.4
    floor = static_rect(0.0, -0.5, 50.0, 0.5, mu_s=0.5, mu_d=mu_d)
    box = body(SHAPE_RECT, 0.0, 0.251, hx=0.5, hy=0.25, vx=6.0, mu_s=0.5, mu_d=mu_d)
    w = world(ref_api, cat(floor, box))
    for _ in range(12):
        w.step()
    v0 = float(w.download_bodies()["vx"][1])
    for _ in range(30):
        w.step()
    v1 = float(w.download_bodies()["vx"][1])
    assert (v0 - v1) / 0.5 == pytest.approx(mu_d * G, rel=0.1)


@pytest.mark.parametrize("order", [_ffi.ORDER_COLOUR, _ffi.ORDER_REFERENCE])
@pytest.mark.parametrize("height", [5, 10])
def test_perturbed_tower_stays_up(ref_api, order, height):
    """a tower of boxes with +-1 cm lateral offsets is statically stable: after 10 s it must
    not have rocked, walked or toppled (this is the test that rejected position-level static
    friction through the contact arms, see solve_contact_manifold)"""
    floor = static_rect(0.0, -0.5, 5.0, 0.5, mu_s=0.6, mu_d=0.4)
    x0 = [0.01 * ((k * 7) % 5 - 2) for k in range(height)]
    boxes = [body(SHAPE_RECT, x0[k], 0.251 + 0.502 * k, hx=0.23, hy=0.23, r=0.02, mu_s=0.6,
                  mu_d=0.4) for k in range(height)]
    w = world(ref_api, cat(floor, *boxes), solver_order=order)
    worst = 0.0
    for _ in range(600):
        w.step()
        worst = max(worst, float(np.abs(w.download_bodies()["angle"][1:]).max()))
    s = w.download_bodies()
    assert worst < 2e-3
    assert np.abs(s["x"][1:] - np.array(x0, dtype=np.float32)).max() < 5e-3
    assert s["y"][height] == pytest.approx(0.25 + 0.5 * (height - 1), abs=5e-3)


def test_box_stack_is_stable(ref_api):
    """5 boxes stacked: after 5 s the top box has not sunk or drifted"""
    floor = static_rect(0.0, -0.5, 5.0, 0.5, mu_s=0.6, mu_d=0.4)
    boxes = [body(SHAPE_RECT, 0.0, 0.251 + 0.502 * k, hx=0.25, hy=0.23, r=0.02, mu_s=0.6, mu_d=0.4)
             for k in range(5)]
    w = world(ref_api, cat(floor, *boxes))
    for _ in range(300):
        w.step()
    s = w.download_bodies()
    top = s[5]
    assert top["y"] == pytest.approx(0.25 + 4 * 0.5, abs=2e-3)
    assert abs(top["x"]) < 1e-3
    assert abs(top["angle"]) < 1e-3
    assert np.abs(s["vy"]).max() < 0.05


def test_contact_event_reports_the_resting_weight(ref_api):
    """a ball resting on the floor: one event per substep-active pair, normal up, point at the
    touching point, impulse = m g h (the weight carried over one substep)"""
    floor = static_rect(0.0, -0.5, 5.0, 0.5)
    ball = body(SHAPE_CIRCLE, 0.0, 0.501, r=0.5)
    w = world(ref_api, cat(floor, ball))
    for _ in range(90):
        w.step()
    ev = w.get_contact_events()
    assert len(ev) == 1 and (ev[0]["a"], ev[0]["b"], ev[0]["points"]) == (0, 1, 1)
    assert (ev[0]["nx"], ev[0]["ny"]) == pytest.approx((0.0, 1.0), abs=1e-6)
    assert (ev[0]["px"], ev[0]["py"]) == pytest.approx((0.0, 0.0), abs=2e-3)
    m = math.pi * 0.25
    assert ev[0]["impulse"] == pytest.approx(m * G / 480.0, rel=0.05)


def test_update_bodies_teleports_only_the_named_bodies(ref_api):
    floor = static_rect(0.0, -0.5, 5.0, 0.5)
    a = body(SHAPE_CIRCLE, -1.0, 0.5, r=0.5)
    b = body(SHAPE_CIRCLE, 1.0, 0.5, r=0.5)
    w = world(ref_api, cat(floor, a, b))
    for _ in range(30):
        w.step()
    before = w.download_bodies()
    st = before[[2]].copy()
    st["x"], st["y"], st["vx"], st["vy"] = 3.0, 4.0, 0.0, 0.0
    w.update_bodies(np.array([2], dtype=np.uint32), st)
    after = w.download_bodies()
    assert after[1].tobytes() == before[1].tobytes()
    assert (after[2]["x"], after[2]["y"]) == (3.0, 4.0)
    w.step()
    assert w.download_bodies()[2]["y"] < 4.0  # it falls from its new place
    with pytest.raises(sf.SfpError):
        w.update_bodies(np.array([7], dtype=np.uint32), st)
