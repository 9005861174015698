"""Seeded sweep over world parameters, sizes, shape mixes, joints and launch modes: the CUDA path
(through the C ABI) against the CPU oracle, bit for bit.

The structured scenes of test_gpu_parity.py use the default parameters; this file moves every
knob the ABI exposes — cell size (down to cells smaller than the bodies and up to one cell for
everything), AABB / contact margins, sweep factor, gravity direction (and none), frame length,
substep count, per-substep broadphase, one-kernel-per-stage mode — over random soups that mix the
four collider kinds with random distance and rope joints between neighbours.  Reference
counterpart: none (the reference tree is not mounted); the bar is the same as everywhere else in
the GPU tier: identical pair set, counts, colours, and identical bits in every body state.
"""
import numpy as np
import pytest

import starframe_b200 as sf
from starframe_b200 import scenes
from starframe_b200._ffi import (FLAG_BROADPHASE_EVERY_SUBSTEP, FLAG_MULTI_KERNEL, JOINT_DESC_DTYPE,
                                 SHAPE_POLYGON)

from test_gpu_parity import assert_same_discrete, assert_same_state

pytestmark = pytest.mark.gpu


def fuzz_scene(n, seed, polygons, joint_fraction):
    b, _ = scenes.random_soup(n, seed, pitch=0.55 + 0.3 * float(scenes.uniform(seed, 40, 1)[0]))
    table = None
    if polygons:
        table = scenes.polygon_table()
        k = np.arange(len(b)) < n  # not the floor
        pick = k & (scenes.uniform(seed, 41, len(b)) < 0.4)
        which = (scenes.uniform(seed, 42, len(b)) * len(table)).astype(np.uint32).clip(0, len(table) - 1)
        b["shape"] = np.where(pick, SHAPE_POLYGON | (which << 8), b["shape"])
        b["radius"] = np.where(pick, np.where(scenes.uniform(seed, 43, len(b)) < 0.5, 0.0, 0.03), b["radius"])
        b["hx"] = np.where(pick, 0.0, b["hx"])
        b["hy"] = np.where(pick, 0.0, b["hy"])
    m = int(joint_fraction * n)
    j = np.zeros(m, dtype=JOINT_DESC_DTYPE)
    if m:
        a = (scenes.uniform(seed, 44, m) * n).astype(np.uint32).clip(0, n - 1)
        # partner: a neighbour in the spawn grid (index +1), wrapped; never the body itself
        bb = (a + 1 + (scenes.uniform(seed, 45, m) * 2).astype(np.uint32)) % max(n, 2)
        bb = np.where(bb == a, (a + 1) % max(n, 2), bb)
        j["body_a"], j["body_b"] = a, bb
        j["ax"] = scenes.uniform(seed, 46, m, -0.05, 0.05)
        j["ay"] = scenes.uniform(seed, 47, m, -0.05, 0.05)
        j["bx"] = scenes.uniform(seed, 48, m, -0.05, 0.05)
        j["by"] = scenes.uniform(seed, 49, m, -0.05, 0.05)
        # rest length near the spawn distance: a joint that starts far from its rest length is
        # resolved within one substep and throws its bodies out of the scene
        d = np.hypot(b["x"][bb] - b["x"][a], b["y"][bb] - b["y"][a])
        j["rest"] = d * scenes.uniform(seed, 50, m, 0.95, 1.1)
        j["compliance"] = np.where(scenes.uniform(seed, 51, m) < 0.5, 0.0, 1.0e-4)
        j["kind"] = (scenes.uniform(seed, 52, m) < 0.5).astype(np.uint32)
        if n < 2:
            j = j[:0]
    return b, j, table


CASES = [
    # n, seed, polygons, joints/body, params, dt, substeps, frames
    (1, 101, False, 0.0, dict(), 1 / 60, 8, 40),
    (2, 102, False, 0.5, dict(cell_size=2.0), 1 / 60, 8, 40),
    (7, 103, True, 0.3, dict(cell_size=0.25, aabb_margin=0.0), 1 / 30, 3, 40),
    (33, 104, False, 0.2, dict(gravity_x=3.0, gravity_y=-4.0, vel_expand=2.0), 1 / 120, 1, 120),
    (257, 105, True, 0.1, dict(cell_size=1.0, contact_margin=0.03), 1 / 60, 16, 40),
    (900, 106, False, 0.25, dict(gravity_x=0.0, gravity_y=0.0), 1 / 60, 8, 30),
    (900, 107, True, 0.0, dict(flags=FLAG_BROADPHASE_EVERY_SUBSTEP), 1 / 60, 4, 50),
    (900, 108, False, 0.2, dict(flags=FLAG_MULTI_KERNEL, cell_size=0.3), 1 / 60, 4, 12),
    (2500, 109, True, 0.15, dict(aabb_margin=0.05, vel_expand=0.5), 1 / 60, 8, 50),
    (2500, 110, False, 0.3, dict(cell_size=4.0, flags=FLAG_BROADPHASE_EVERY_SUBSTEP), 1 / 90, 2, 80),
    (1500, 111, True, 0.2, dict(cell_size=0.15, max_entries=400000), 1 / 60, 8, 40),
    (4000, 112, False, 0.05, dict(gravity_y=-30.0), 1 / 60, 8, 40),
]


@pytest.mark.parametrize("case", CASES, ids=[f"n{c[0]}-s{c[1]}" for c in CASES])
def test_parameter_sweep_matches_oracle(gpu_api, ref_api, case):
    n, seed, polygons, jf, params, dt, substeps, frames = case
    b, j, table = fuzz_scene(n, seed, polygons, jf)
    worlds = []
    for api in (gpu_api, ref_api):
        w = sf.World(api, **params)
        if table is not None:
            w.set_polygons(table)
        w.upload_bodies(b)
        w.set_joints(j)
        worlds.append(w)
    g, r = worlds
    for f in range(frames):
        g.step(dt, substeps)
        r.step(dt, substeps)
        if f % 8 == 7 or f == frames - 1:
            assert_same_discrete(g, r)
            assert_same_state(g, r)
    # the sweep is only worth something if the worlds are busy
    if n >= 257 and frames >= 40 and params.get("gravity_y", -9.81) != 0.0:
        assert r.stats()["n_contacts"] > n // 8


def test_reupload_and_joint_changes_mid_run(gpu_api, ref_api):
    """bodies re-uploaded (different count), joints replaced and removed between frames: the
    warm-started colour table and the captured launch graphs must not leak across uploads"""
    g, r = sf.World(gpu_api), sf.World(ref_api)
    for k, (n, seed, jf) in enumerate([(600, 201, 0.2), (350, 202, 0.0), (900, 203, 0.3)]):
        b, j, _ = fuzz_scene(n, seed, False, jf)
        for w in (g, r):
            w.upload_bodies(b)
            w.set_joints(j)
        for f in range(5):
            g.step()
            r.step()
        assert_same_discrete(g, r)
        assert_same_state(g, r)
        if k == 0:
            # drop the joints, keep the bodies
            for w in (g, r):
                w.set_joints(j[:0])
            for f in range(3):
                g.step()
                r.step()
            assert_same_discrete(g, r)
            assert_same_state(g, r)
