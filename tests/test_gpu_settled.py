"""The regime bench.py times — the SETTLED scene, contacts dominant, full size — against the
oracle, and the same with the bodies uploaded in a random order.

Each scene is settled for 240 frames on the GPU, the state is downloaded and handed to a fresh GPU
world and a fresh oracle world (same descs, same state through sfp_step_host), both step two
frames; pairs, counts, colours and every state bit must agree.  (The oracle does a settled
100 k-body frame in ~0.5 s; settling it there would take minutes.)

Shuffled uploads: the library keeps its bodies in an order of its own behind a permutation
(DESIGN.md §3); everything whose result depends on body numbering uses the caller's indices, so
GPU(shuffled) == oracle(shuffled) bit for bit as well.  GPU(shuffled) == GPU(unshuffled) holds for
what does not depend on numbering — the candidate pair SET exactly, the contact count of the
first substep up to points sitting at the contact margin — but not bit for bit for the state: which body of a pair is "a" and the colouring
priorities follow the caller's indices, i.e. the Gauss-Seidel order is a different (equally valid)
one.
"""
import os

import numpy as np
import pytest

import starframe_b200 as sf
from starframe_b200 import scenes

pytestmark = pytest.mark.gpu
FIELDS = ("x", "y", "c", "s", "vx", "vy", "omega", "angle")
SETTLE = 240


def shuffled(bodies, joints, seed=20261107):
    perm = np.random.default_rng(seed).permutation(len(bodies))  # new index i holds old body perm[i]
    inv = np.empty_like(perm)
    inv[perm] = np.arange(len(bodies))
    j = joints.copy()
    if len(j):
        j["body_a"] = inv[joints["body_a"]]
        j["body_b"] = inv[joints["body_b"]]
    return bodies[perm], j, perm, inv


def oracle_world(ref_api, **params):
    old = os.environ.get("SFP_REF_THREADS")
    os.environ["SFP_REF_THREADS"] = str(min(os.cpu_count() or 1, 32))
    try:
        return sf.World(ref_api, **params)
    finally:
        if old is None:
            os.environ.pop("SFP_REF_THREADS", None)
        else:
            os.environ["SFP_REF_THREADS"] = old


def settle_then_compare(gpu_api, ref_api, b, j, frames=2, **params):
    w = sf.World(gpu_api, **params)
    w.upload_bodies(b)
    w.set_joints(j)
    for _ in range(SETTLE):
        w.step()
    settled = w.download_bodies()
    st0 = w.stats()
    assert st0["overflow"] == 0
    del w
    g, r = sf.World(gpu_api, **params), oracle_world(ref_api, **params)
    for x in (g, r):
        x.upload_bodies(b)
        x.set_joints(j)
    sg = sr = settled
    for _ in range(frames):
        sg = g.step_host(sg)
        sr = r.step_host(sr)
    pg, pr = g.get_pairs(), r.get_pairs()
    assert pg.shape == pr.shape and np.array_equal(pg, pr)
    a, c = g.stats(), r.stats()
    for k in ("n_entries", "n_pairs", "n_contacts", "n_manifolds", "n_colours", "colour_rounds"):
        assert a[k] == c[k], (k, a[k], c[k])
    assert np.float32(a["max_penetration"]) == np.float32(c["max_penetration"])
    cg, cr = g.get_colours(), r.get_colours()
    assert np.array_equal(cg[1], cr[1]) and np.array_equal(cg[2], cr[2])
    for f in FIELDS:
        bad = np.nonzero(sg[f].view(np.uint32) != sr[f].view(np.uint32))[0]
        assert bad.size == 0, (f, bad[:5], sg[f][bad[:5]], sr[f][bad[:5]])
    # the regime: the settled scene is contact dominated
    return a


@pytest.mark.parametrize("name", ["pile100k", "ropes2k", "circles10k"])
def test_settled_scene_matches_oracle(gpu_api, ref_api, name):
    from bench import make_scene

    b, j = make_scene(name)
    st = settle_then_compare(gpu_api, ref_api, b, j)
    if name == "pile100k":
        assert st["n_pairs"] > 3.0 * len(b) and st["n_contacts"] > 2.5 * len(b)


@pytest.mark.parametrize("name", ["pile100k", "ropes2k", "circles10k"])
def test_settled_scene_shuffled_upload_matches_oracle(gpu_api, ref_api, name):
    from bench import make_scene

    b, j = make_scene(name)
    bs, js, perm, inv = shuffled(b, j)
    settle_then_compare(gpu_api, ref_api, bs, js)


def test_pair_set_and_first_contacts_do_not_depend_on_upload_order(gpu_api):
    b, j = scenes.mixed_pile(20_000)
    # start from an overlapping state so that the first substep already has contacts
    w0 = sf.World(gpu_api)
    w0.upload_bodies(b)
    for _ in range(200):
        w0.step()
    settled = w0.download_bodies()
    bs, js, perm, inv = shuffled(b, j)
    res = []
    for bodies, state in ((b, settled), (bs, settled[perm])):
        w = sf.World(gpu_api)
        w.upload_bodies(bodies)
        w.step_host(state, 1.0 / 480.0, 1)
        res.append((w.get_pairs(), w.stats()))
    (p0, s0), (p1, s1) = res
    # pairs of the shuffled world in the unshuffled numbering
    a, c = perm[p1["a"]], perm[p1["b"]]
    lo, hi = np.minimum(a, c), np.maximum(a, c)
    k1 = np.sort(lo.astype(np.int64) << 32 | hi.astype(np.int64))
    k0 = np.sort(p0["a"].astype(np.int64) << 32 | p0["b"].astype(np.int64))
    assert np.array_equal(k0, k1)
    # which body of a pair is "a" follows the caller's numbering and the narrowphase is symmetric
    # under that swap only up to rounding: a handful of contact points sitting exactly at the
    # contact margin come and go (measured: 5 of 57 169)
    assert abs(int(s0["n_contacts"]) - int(s1["n_contacts"])) <= 1e-3 * s0["n_contacts"]
    assert abs(int(s0["n_manifolds"]) - int(s1["n_manifolds"])) <= 1e-3 * s0["n_manifolds"]


def test_settled_batch_matches_oracle(gpu_api, ref_api):
    """one GPU's share of config 5 (1024 worlds x 512 bodies), settled, then two frames through
    sfp_batch_step_host against oracle worlds fed the same states (a sample of 48 worlds)"""
    nw, bpw = 1024, 512
    bodies = scenes.worlds(nw, bpw)
    batch = sf.Batch(gpu_api, nw, bpw)
    batch.upload(bodies)
    for _ in range(SETTLE):
        batch.step()
    settled = batch.download()
    assert (batch.stats()["overflow"] == 0).all()
    del batch
    g = sf.Batch(gpu_api, nw, bpw)
    g.upload(bodies)
    sg = settled.reshape(-1)
    for _ in range(2):
        sg = g.step_host(sg).reshape(-1)
    sg = sg.reshape(nw, bpw)
    stats = g.stats()
    for wi in list(range(0, nw, 23)) + [nw - 1]:
        r = sf.World(ref_api)
        r.upload_bodies(bodies[wi])
        sr = settled[wi]
        for _ in range(2):
            sr = r.step_host(sr)
        for f in FIELDS:
            assert np.array_equal(sg[wi][f].view(np.uint32), sr[f].view(np.uint32)), (wi, f)
        rs = r.stats()
        assert stats["n_pairs"][wi] == rs["n_pairs"] and stats["n_contacts"][wi] == rs["n_contacts"]


@pytest.mark.parametrize("scene", ["pile", "ropes", "polys", "plank"])
def test_tile_order_matches_oracle(gpu_api, ref_api, scene):
    """the optional tile solve order (SFP_FLAG_TILES, sfp_math.h "Tiles"): GPU vs oracle, incl.
    the re-sort of frame 32 and a shuffled upload"""
    from starframe_b200._ffi import FLAG_TILES

    polys = None
    if scene == "pile":
        b, j = scenes.mixed_pile(3000)
    elif scene == "ropes":
        b, j = scenes.ropes(n_ropes=40, links=32, n_free=800)
    elif scene == "polys":
        b, j, polys = scenes.polygon_pile(1500)
    else:
        b, j = scenes.crowded_plank(100)
    for do_shuffle in (False, True):
        bb, jj = (shuffled(b, j)[:2] if do_shuffle else (b, j))
        ws = []
        for api in (gpu_api, ref_api):
            w = sf.World(api, flags=FLAG_TILES)
            if polys is not None:
                w.set_polygons(polys)
            w.upload_bodies(bb)
            w.set_joints(jj)
            ws.append(w)
        g, r = ws
        for f in range(45):
            g.step()
            r.step()
            if f % 11 == 10 or f == 44:
                assert np.array_equal(g.get_pairs(), r.get_pairs())
                cg, cr = g.get_colours(), r.get_colours()
                assert np.array_equal(cg[1], cr[1]) and np.array_equal(cg[2], cr[2])
                a, c = g.download_bodies(), r.download_bodies()
                for fld in FIELDS:
                    assert np.array_equal(a[fld].view(np.uint32), c[fld].view(np.uint32)), (scene, f, fld)
        if scene == "pile" and not do_shuffle:
            assert np.mean(g.get_colours()[1] >= 64) > 0.8  # most constraints are tile-interior
