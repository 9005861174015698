"""GPU (libsfp_b200.so, through the C ABI) vs the CPU oracle on identical input state.

Bars (BASELINE.json north_star): sorted candidate-pair set and contact counts BIT-EXACT;
positions / velocities within a stated tolerance.  The oracle in SFP_ORDER_COLOUR performs the
same Gauss-Seidel sweep in colour order with the same arithmetic (include/sfp_math.h, FMA
contraction off on both sides), so the tolerance used here is 0: bit-exact state.  The looser
1e-4 relative bar that accounts for Gauss-Seidel ORDER (reference order vs colour order) is
measured oracle-vs-oracle in test_oracle_props.py.
"""
import numpy as np
import pytest

import starframe_b200 as sf
from starframe_b200 import scenes

pytestmark = pytest.mark.gpu

FIELDS = ("x", "y", "c", "s", "vx", "vy", "omega")


def make_worlds(gpu_api, ref_api, bodies, joints, **params):
    g = sf.World(gpu_api, **params)
    r = sf.World(ref_api, **params)
    for w in (g, r):
        w.upload_bodies(bodies)
        w.set_joints(joints)
    return g, r


def assert_same_state(g, r, exact=True, tol=0.0):
    a, b = g.download_bodies(), r.download_bodies()
    for f in FIELDS:
        if exact:
            bad = np.nonzero(a[f].view(np.uint32) != b[f].view(np.uint32))[0]
            assert bad.size == 0, (f, bad[:8], a[f][bad[:8]], b[f][bad[:8]])
        else:
            err = np.max(np.abs(a[f] - b[f]) / (1.0 + np.abs(b[f])))
            assert err <= tol, (f, err)


def assert_same_discrete(g, r):
    pg, pr = g.get_pairs(), r.get_pairs()
    assert pg.shape == pr.shape, (pg.shape, pr.shape)
    assert np.array_equal(pg, pr)
    sg, sr = g.stats(), r.stats()
    for k in ("n_entries", "n_pairs", "n_contacts", "n_manifolds", "n_colours", "colour_rounds"):
        assert sg[k] == sr[k], (k, sg[k], sr[k])
    assert np.float32(sg["max_penetration"]) == np.float32(sr["max_penetration"])
    _, cg_, jg = g.get_colours()
    _, cr_, jr = r.get_colours()
    assert np.array_equal(cg_, cr_)
    assert np.array_equal(jg, jr)


def run_and_compare(g, r, frames, dt=1 / 60, substeps=8, every=1):
    for f in range(frames):
        g.step(dt, substeps)
        r.step(dt, substeps)
        if (f + 1) % every == 0 or f == frames - 1:
            assert_same_discrete(g, r)
            assert_same_state(g, r)


def test_single_substep_stack(gpu_api, ref_api):
    b, j = scenes.stack100()
    g, r = make_worlds(gpu_api, ref_api, b, j)
    g.step(1 / 480, 1)
    r.step(1 / 480, 1)
    assert_same_discrete(g, r)
    assert_same_state(g, r)


def test_stack100_rollout(gpu_api, ref_api):
    b, j = scenes.stack100()
    g, r = make_worlds(gpu_api, ref_api, b, j)
    run_and_compare(g, r, 120, every=10)
    cg_, cr_ = g.get_contacts(), r.get_contacts()
    assert cg_.shape == cr_.shape
    assert cg_.tobytes() == cr_.tobytes()


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_random_soup(gpu_api, ref_api, seed):
    b, j = scenes.random_soup(600, seed)
    g, r = make_worlds(gpu_api, ref_api, b, j)
    run_and_compare(g, r, 20, every=5)


def test_multi_kernel_mode_matches(gpu_api, ref_api):
    b, j = scenes.random_soup(400, 11)
    g, r = make_worlds(gpu_api, ref_api, b, j, flags=sf._ffi.FLAG_MULTI_KERNEL)
    r2 = sf.World(ref_api)
    run_and_compare(g, r, 5)


@pytest.mark.parametrize("scene", ["pile", "ropes"])
def test_across_recolourings_and_resorts(gpu_api, ref_api, scene):
    """100 frames: every 32nd frame (SFP_RECOLOUR_PERIOD) both sides colour from scratch and the
    device renews its body order; compared around those frames and at the end, bodies uploaded in
    a random order so that the re-sorts really move them"""
    if scene == "pile":
        b, j = scenes.mixed_pile(700)
    else:
        b, j = scenes.ropes(n_ropes=16, links=20, n_free=200)
    perm = np.random.default_rng(5).permutation(len(b))
    inv = np.empty_like(perm)
    inv[perm] = np.arange(len(b))
    b = b[perm]
    if len(j):
        j = j.copy()
        j["body_a"], j["body_b"] = inv[j["body_a"]], inv[j["body_b"]]
    g, r = make_worlds(gpu_api, ref_api, b, j)
    for f in range(100):
        g.step()
        r.step()
        if f in (30, 31, 32, 33, 63, 64, 65, 96, 99):
            assert_same_discrete(g, r)
            assert_same_state(g, r)
    assert g.stats()["n_contacts"] > 300


def test_circles_container_small(gpu_api, ref_api):
    b, j = scenes.circles_container(2000)
    g, r = make_worlds(gpu_api, ref_api, b, j)
    run_and_compare(g, r, 60, every=20)


def test_mixed_pile_small(gpu_api, ref_api):
    b, j = scenes.mixed_pile(3000)
    g, r = make_worlds(gpu_api, ref_api, b, j)
    run_and_compare(g, r, 60, every=20)


def test_ropes_small(gpu_api, ref_api):
    b, j = scenes.ropes(n_ropes=40, links=32, n_free=300)
    g, r = make_worlds(gpu_api, ref_api, b, j)
    run_and_compare(g, r, 40, every=10)


@pytest.mark.parametrize("flags", [0, sf._ffi.FLAG_MULTI_KERNEL])
def test_serial_group_plank(gpu_api, ref_api, flags):
    """a body with more partners than colours: the serial group (colour 63) on the device"""
    b, j = scenes.crowded_plank(100)
    g, r = make_worlds(gpu_api, ref_api, b, j, flags=flags)
    run_and_compare(g, r, 40, every=10)
    assert g.stats()["n_colours"] == 64


def test_contact_events_match_oracle(gpu_api, ref_api):
    """device-compacted contact events == the oracle's, field for field"""
    b, j = scenes.mixed_pile(2500)
    g, r = make_worlds(gpu_api, ref_api, b, j)
    assert len(g.get_contact_events()) == 0
    for f in range(120):
        g.step()
        r.step()
        if f % 40 == 39:
            eg, er = g.get_contact_events(), r.get_contact_events()
            assert len(eg) == len(er) and len(eg) > 0
            assert eg.tobytes() == er.tobytes()
    assert len(eg) > 1000


def test_update_bodies_matches_oracle(gpu_api, ref_api):
    b, j = scenes.mixed_pile(1500)
    g, r = make_worlds(gpu_api, ref_api, b, j)
    rng = np.random.default_rng(3)
    for round_ in range(4):
        for _ in range(15):
            g.step()
            r.step()
        idx = rng.choice(np.arange(3, len(b)), size=40, replace=False).astype(np.uint32)
        st = r.download_bodies()[idx].copy()
        st["y"] += 3.0
        st["vx"] = rng.uniform(-2, 2, 40).astype(np.float32)
        if round_ == 1:
            st["c"] = 0.0
            st["s"] = 0.0  # rotation given as an angle only
        g.update_bodies(idx, st)
        r.update_bodies(idx, st)
    for _ in range(10):
        g.step()
        r.step()
    assert_same_discrete(g, r)
    assert_same_state(g, r)


def test_step_host_roundtrip(gpu_api, ref_api):
    b, j = scenes.mixed_pile(1500)
    g, r = make_worlds(gpu_api, ref_api, b, j)
    sg, sr = g.download_bodies(), r.download_bodies()
    for _ in range(5):
        sg = g.step_host(sg)
        sr = r.step_host(sr)
    for f in FIELDS:
        assert np.array_equal(sg[f].view(np.uint32), sr[f].view(np.uint32)), f


def test_step_host_with_pinned_arrays(gpu_api, ref_api):
    """states kept in sfp_alloc_host_states memory take the direct-DMA path: same result"""
    b, j = scenes.mixed_pile(1500)
    g, r = make_worlds(gpu_api, ref_api, b, j)
    n = len(b)
    buf = [gpu_api.alloc_host_states(n), gpu_api.alloc_host_states(n)]
    buf[0][:] = g.download_bodies()
    sr = r.download_bodies()
    for k in range(6):
        g.step_host(buf[k & 1], out=buf[(k + 1) & 1])
        sr = r.step_host(sr)
    for f in FIELDS:
        assert np.array_equal(buf[0][f].view(np.uint32), sr[f].view(np.uint32)), f


@pytest.mark.parametrize("pinned", [False, True])
def test_states_with_zero_rotor_take_their_angle(gpu_api, ref_api, pinned):
    """a scene graph that keeps angles only hands over states with c = s = 0: both sides build
    the rotor from `angle` with the shared sincos_det (no libm / CUDA sincosf involved)"""
    b, j = scenes.mixed_pile(800)
    g, r = make_worlds(gpu_api, ref_api, b, j)
    n = len(b)
    sg = gpu_api.alloc_host_states(n) if pinned else np.zeros(n, dtype=sf._ffi.BODY_STATE_DTYPE)
    out = gpu_api.alloc_host_states(n) if pinned else np.zeros(n, dtype=sf._ffi.BODY_STATE_DTYPE)
    sg[:] = g.download_bodies()
    sr = r.download_bodies()
    for k in range(4):
        for s_ in (sg, sr):
            s_["angle"][::3] += np.float32(0.25 * (k + 1))  # the game turns every third body ...
            s_["c"][::3] = 0.0                               # ... and only knows angles
            s_["s"][::3] = 0.0
        g.step_host(sg, out=out)
        sr = r.step_host(sr)
        sg[:] = out
    for f in FIELDS + ("angle",):
        assert np.array_equal(sg[f].view(np.uint32), sr[f].view(np.uint32)), f
    # and update_bodies takes the same route
    idx = np.arange(0, n, 7, dtype=np.uint32)
    st = sr[idx].copy()
    st["c"], st["s"] = 0.0, 0.0
    st["angle"] = np.linspace(-9.0, 9.0, len(idx)).astype(np.float32)
    g.update_bodies(idx, st)
    r.update_bodies(idx, st)
    assert_same_state(g, r)
    got = r.download_bodies()[idx]
    assert np.max(np.abs(got["c"] - np.cos(st["angle"].astype(np.float64)))) < 3e-7
    assert np.max(np.abs(got["s"] - np.sin(st["angle"].astype(np.float64)))) < 3e-7


def test_empty_and_tiny_worlds(gpu_api, ref_api):
    for n in (0, 1, 2):
        b, j = scenes.random_soup(max(n, 1), 5)
        b = b[:n]
        g, r = make_worlds(gpu_api, ref_api, b, j)
        g.step()
        r.step()
        g.sync()
        assert g.stats()["n_pairs"] == r.stats()["n_pairs"]
        if n:
            assert_same_state(g, r)


def test_capacity_overflow_reported(gpu_api):
    b, j = scenes.random_soup(500, 9, extent=1.0)
    g = sf.World(gpu_api, max_pairs=16)
    g.upload_bodies(b)
    g.step()
    with pytest.raises(sf.SfpError) as e:
        g.sync()
    assert e.value.code == sf._ffi.SFP_E_CAPACITY


def test_determinism(gpu_api):
    b, j = scenes.mixed_pile(2000)
    outs = []
    for _ in range(2):
        g = sf.World(gpu_api)
        g.upload_bodies(b)
        for _ in range(30):
            g.step()
        outs.append(g.download_bodies().tobytes())
    assert outs[0] == outs[1]


def test_worlds_driven_from_different_threads(gpu_api, ref_api):
    """one world per thread at a time, different worlds from different threads.
    Four threads step four worlds (cooperative launches, graphs and all) at the same time; every
    world must still equal its oracle twin bit for bit."""
    import threading

    seeds = [21, 22, 23, 24]
    worlds, results, errors = [], [None] * len(seeds), []
    for sd in seeds:
        b, j = scenes.random_soup(700, sd)
        g, r = make_worlds(gpu_api, ref_api, b, j)
        worlds.append((g, r))

    def run(k):
        try:
            g = worlds[k][0]
            for _ in range(25):
                g.step()
            results[k] = g.download_bodies()
        except Exception as e:  # surfaced below: an assertion in a thread is otherwise silent
            errors.append((k, e))

    threads = [threading.Thread(target=run, args=(k,)) for k in range(len(seeds))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for k, (g, r) in enumerate(worlds):
        for _ in range(25):
            r.step()
        ro = r.download_bodies()
        for f in FIELDS:
            assert np.array_equal(results[k][f].view(np.uint32), ro[f].view(np.uint32)), (k, f)


@pytest.mark.parametrize("cell,frames_queued", [(0.5, 1), (0.1, 3)])
def test_capacity_overflow_grows_and_redoes_the_frames(gpu_api, ref_api, cell, frames_queued):
    """A dense blob has several times more candidate pairs (and, with small cells, entries) than
    the default capacities (12 n + 4096 pairs, 8 n + 8192 entries) hold.  The library notices at
    the next synchronisation, enlarges its buffers, puts the pre-frame state back and runs the
    queued frames again: the result must be the oracle's (which has no capacities) bit for bit.
    With explicit max_pairs the same world is an SFP_E_CAPACITY error."""
    from starframe_b200._ffi import SfpError, SFP_E_CAPACITY

    # a loose grid of circles with a wide AABB margin: ~22 candidate pairs per body (few of them
    # touch), and with 0.1 m cells ~70 broadphase entries per body
    n_side, pitch = 45, 0.2
    b = scenes.empty_bodies(n_side * n_side)
    k = np.arange(n_side * n_side)
    b["x"] = pitch * (k % n_side) + 0.003 * scenes.uniform(99, 1, len(k))
    b["y"] = 0.12 + pitch * (k // n_side) + 0.003 * scenes.uniform(99, 2, len(k))
    b["vy"] = -1.0
    b["shape"], b["radius"] = 0, 0.09
    b = np.concatenate([b, scenes.static_rect(4.4, -0.5, 8.0, 0.5)])
    j = scenes.no_joints()
    n = len(b)
    g, r = make_worlds(gpu_api, ref_api, b, j, cell_size=cell, aabb_margin=0.26)
    for _ in range(2):
        for _ in range(frames_queued):
            g.step()
            r.step()
        assert_same_discrete(g, r)   # (the first call that synchronises triggers the recovery)
        assert_same_state(g, r)
    st = g.stats()
    assert st["n_pairs"] > 12 * n + 4096, st["n_pairs"]
    if cell < 0.2:
        assert st["n_entries"] > 8 * n + 8192, st["n_entries"]
    assert st["overflow"] == 0
    # explicit capacity: a contract, not a hint
    fixed = sf.World(gpu_api, cell_size=cell, aabb_margin=0.26, max_pairs=12 * n + 4096)
    fixed.upload_bodies(b)
    fixed.step()
    with pytest.raises(SfpError) as e:
        fixed.sync()
    assert e.value.code == SFP_E_CAPACITY
