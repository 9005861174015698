"""Batched independent worlds (BASELINE.json config 5) on the GPU, through the C ABI:
world w stepped inside a batch == world w stepped alone on the GPU == the CPU oracle, bit for bit."""
import numpy as np
import pytest

import starframe_b200 as sf
from starframe_b200 import scenes

pytestmark = pytest.mark.gpu
FIELDS = ("x", "y", "c", "s", "vx", "vy", "omega")


def same(a, b):
    for f in FIELDS:
        bad = np.nonzero(a[f].view(np.uint32) != b[f].view(np.uint32))[0]
        assert bad.size == 0, (f, bad[:5], a[f][bad[:5]], b[f][bad[:5]])


@pytest.mark.parametrize("bpw,nw,frames", [(128, 6, 60), (512, 3, 90)])
def test_batch_equals_oracle_and_single_world(gpu_api, ref_api, bpw, nw, frames):
    bodies = scenes.worlds(nw, bpw)
    batch = sf.Batch(gpu_api, nw, bpw)
    batch.upload(bodies)
    refs, singles = [], []
    for w in range(nw):
        r = sf.World(ref_api)
        r.upload_bodies(bodies[w])
        refs.append(r)
    g = sf.World(gpu_api)
    g.upload_bodies(bodies[nw - 1])
    for f in range(frames):
        batch.step()
        for r in refs:
            r.step()
        g.step()
    out = batch.download()
    st = batch.stats()
    for w in range(nw):
        same(out[w], refs[w].download_bodies())
        rs = refs[w].stats()
        for k in ("n_pairs", "n_contacts", "n_manifolds", "n_colours", "colour_rounds"):
            assert st[k][w] == rs[k], (w, k, st[k][w], rs[k])
        assert np.float32(st["max_penetration"][w]) == np.float32(rs["max_penetration"])
    same(out[nw - 1], g.download_bodies())
    assert st["n_contacts"].sum() > 0


def test_batch_many_worlds_spot_check(gpu_api, ref_api):
    """more worlds than resident CTAs: every world must still be stepped exactly once per frame"""
    nw, bpw = 700, 64
    bodies = scenes.worlds(nw, bpw)
    batch = sf.Batch(gpu_api, nw, bpw)
    batch.upload(bodies)
    for _ in range(40):
        batch.step()
    out = batch.download()
    for w in (0, 1, 299, 444, 699):
        r = sf.World(ref_api)
        r.upload_bodies(bodies[w])
        for _ in range(40):
            r.step()
        same(out[w], r.download_bodies())


@pytest.mark.parametrize("nw,bpw,substeps", [(700, 64, 8), (1000, 48, 3), (445, 64, 1), (2100, 32, 5)])
def test_batch_queue_schedule_equals_one_cta_per_world(gpu_api, nw, bpw, substeps):
    """More worlds than resident CTAs: the frame kernel's CTAs take worlds from a queue, the last
    world per CTA one stage at a time with the state handed from CTA to CTA through
    global memory.  EVERY world must end in the bits it gets with one CTA per world
    (SFP_FLAG_BATCH_WHOLE_WORLDS), state and statistics, after frames enough for contacts, warm
    colourings and recolourings."""
    bodies = scenes.worlds(nw, bpw)
    outs, stats = [], []
    for flags in (0, sf.FLAG_BATCH_WHOLE_WORLDS):
        batch = sf.Batch(gpu_api, nw, bpw, flags=flags)
        batch.upload(bodies)
        for _ in range(24):
            batch.step(substeps=substeps)
        outs.append(batch.download())
        stats.append(batch.stats())
        del batch
    same(outs[0].reshape(-1), outs[1].reshape(-1))
    for k in ("n_pairs", "n_contacts", "n_manifolds", "n_colours", "colour_rounds", "overflow"):
        assert np.array_equal(stats[0][k], stats[1][k]), k
    assert stats[0]["n_contacts"].sum() > 0


@pytest.mark.parametrize("nw", [1200, 2100])
def test_batch_queue_sliced_host_path_matches_whole_worlds(gpu_api, nw):
    """the sliced sfp_batch_step_host path (four slices, the outer two small; the last slice goes
    through the queue by the stage)"""
    bpw, frames = 32, 12
    bodies = scenes.worlds(nw, bpw)
    outs = []
    for flags in (0, sf.FLAG_BATCH_WHOLE_WORLDS):
        batch = sf.Batch(gpu_api, nw, bpw, flags=flags)
        batch.upload(bodies)
        bufs = [gpu_api.alloc_host_states(nw * bpw), gpu_api.alloc_host_states(nw * bpw)]
        bufs[0][:] = batch.download().reshape(-1)
        for f in range(frames):
            batch.step_host(bufs[f & 1], out=bufs[(f + 1) & 1])
        outs.append(np.array(bufs[frames & 1]))
        del batch
    same(outs[0], outs[1])
    assert np.array_equal(outs[0]["angle"].view(np.uint32), outs[1]["angle"].view(np.uint32))


def test_batch_broadphase_sweep_with_many_wide_bodies(gpu_api, ref_api):
    """the per-world broadphase sorts the boxes along x and keeps bodies wider than 8 cells in a list
    of 32 that is tested against everybody: worlds with MORE than 32 such bodies (long bars, some
    static, some falling), an odd body count, and bodies sharing their low edge exactly"""
    nw = 5
    worlds = []
    for w in range(nw):
        b, _ = scenes.random_soup(130, 300 + w, extent=5.0)   # 130 + the floor: an odd count
        rng = np.random.default_rng(w)
        bars = np.arange(0, 44)                      # 44 bars, 6 to 9 m long: all "wide"
        b["shape"][bars] = sf._ffi.SHAPE_RECT
        b["hx"][bars] = rng.uniform(3.0, 4.5, len(bars))
        b["hy"][bars] = 0.05
        b["radius"][bars] = 0.0
        b["angle"][bars] = rng.uniform(-0.2, 0.2, len(bars))
        b["y"][bars] = 0.5 + 0.35 * bars
        b["x"][bars] = rng.uniform(-1.0, 1.0, len(bars))
        b["flags"][bars[::3]] |= sf._ffi.BODY_STATIC
        same_edge = np.arange(60, 80)                # 20 circles with the same low edge
        b["shape"][same_edge] = sf._ffi.SHAPE_CIRCLE
        b["hx"][same_edge] = 0.0
        b["hy"][same_edge] = 0.0
        b["radius"][same_edge] = 0.1
        b["x"][same_edge] = 2.0
        b["vx"][same_edge] = 0.0
        b["omega"][same_edge] = 0.0
        worlds.append(b)
    bodies = np.stack(worlds)
    n = bodies.shape[1]
    batch = sf.Batch(gpu_api, nw, n)
    batch.upload(bodies)
    refs = []
    for w in range(nw):
        r = sf.World(ref_api)
        r.upload_bodies(bodies[w])
        refs.append(r)
    for f in range(25):
        batch.step()
        for r in refs:
            r.step()
    out, st = batch.download(), batch.stats()
    for w in range(nw):
        rs = refs[w].stats()
        for k in ("n_pairs", "n_contacts", "n_colours"):
            assert st[k][w] == rs[k], (w, k, st[k][w], rs[k])
        same(out[w], refs[w].download_bodies())
    assert st["n_pairs"].min() > 150


def test_batch_partition_is_world_index_independent(gpu_api):
    """a rank that owns worlds [k, k+m) computes what a rank owning [0, n) computes for them
    (this is the multi-GPU sharding: no data-path exchange, so no collective to test)"""
    bpw = 128
    bodies = scenes.worlds(8, bpw)
    full = sf.Batch(gpu_api, 8, bpw)
    full.upload(bodies)
    part = sf.Batch(gpu_api, 3, bpw)
    part.upload(scenes.worlds(3, bpw, first_world=4))
    for _ in range(30):
        full.step()
        part.step()
    a, b = full.download(), part.download()
    for k in range(3):
        same(a[4 + k], b[k])


def test_small_batch_created_after_a_large_one_does_not_break_it(gpu_api):
    """the kernel's shared-memory limit is per function, not per batch: creating a batch of
    small worlds must not lower it under a live batch of larger worlds"""
    big = sf.Batch(gpu_api, 2, 512)
    big.upload(scenes.worlds(2, 512))
    big.step()
    small = sf.Batch(gpu_api, 2, 32)
    small.upload(scenes.worlds(2, 32))
    small.step()
    big.step()
    big.sync()
    small.sync()


def test_batch_serial_group(gpu_api, ref_api):
    b, _ = scenes.crowded_plank(90)
    n = len(b)
    batch = sf.Batch(gpu_api, 2, n)
    batch.upload(np.stack([b, b]))
    r = sf.World(ref_api)
    r.upload_bodies(b)
    for _ in range(40):
        batch.step()
        r.step()
    out = batch.download()
    same(out[0], r.download_bodies())
    same(out[1], r.download_bodies())
    assert batch.stats()["n_colours"][0] == 64


def test_batch_rejects_bad_arguments(gpu_api):
    with pytest.raises(sf.SfpError):
        sf.Batch(gpu_api, 0, 64)
    with pytest.raises(sf.SfpError):
        sf.Batch(gpu_api, 4, 0)
    b = sf.Batch(gpu_api, 2, 16)
    with pytest.raises(sf.SfpError):
        b.step()  # nothing uploaded


@pytest.mark.parametrize("pinned,nw,bpw", [(False, 5, 128), (True, 5, 128), (True, 70, 64)])
def test_batch_step_host_matches_oracle(gpu_api, ref_api, pinned, nw, bpw):
    """sfp_batch_step_host: states in from the host every frame (edited by the "game" in between),
    one frame, states out — against the oracle's sfp_ref_step_host per world, bit for bit
    (64 worlds and more with page-locked arrays take the sliced, overlapped path)"""
    frames = 30
    bodies = scenes.worlds(nw, bpw)
    batch = sf.Batch(gpu_api, nw, bpw)
    batch.upload(bodies)
    refs = []
    for w in range(nw):
        r = sf.World(ref_api)
        r.upload_bodies(bodies[w])
        refs.append(r)
    n = nw * bpw
    if pinned:
        bufs = [gpu_api.alloc_host_states(n), gpu_api.alloc_host_states(n)]
    else:
        bufs = [np.zeros(n, dtype=sf._ffi.BODY_STATE_DTYPE) for _ in range(2)]
    bufs[0][:] = batch.download().reshape(-1)
    ref_states = [r.download_bodies() for r in refs]
    for f in range(frames):
        src, dst = bufs[f & 1], bufs[(f + 1) & 1]
        if f == 10:
            # the game kicks one dynamic body of every world (body 5 is dynamic: 0..2 are the container)
            for w in range(nw):
                src[w * bpw + 5]["vx"] = 2.0
                src[w * bpw + 5]["vy"] = 3.0
                ref_states[w][5]["vx"] = 2.0
                ref_states[w][5]["vy"] = 3.0
        batch.step_host(src, out=dst)
        for w in range(nw):
            ref_states[w] = refs[w].step_host(ref_states[w])
    out = bufs[frames & 1].reshape(nw, bpw)
    for w in range(nw):
        same(out[w], ref_states[w])
        assert np.array_equal(out[w]["angle"].view(np.uint32), ref_states[w]["angle"].view(np.uint32))
    assert batch.stats()["n_contacts"].sum() > 0


def rope_worlds(nw):
    """nw small rope worlds of identical size but different content; world w keeps only its first
    (jmax - w) joints, so the per-world counts differ"""
    worlds, joints = [], []
    for w in range(nw):
        b, j = scenes.ropes(n_ropes=4, links=12, n_free=40 + 0 * w, seed=scenes.BASE_SEED + 40 + w, rope_rows=2)
        worlds.append(b)
        joints.append(j)
    n = len(worlds[0])
    assert all(len(b) == n for b in worlds)
    jmax = max(len(j) for j in joints)
    counts = np.array([len(j) - w for w, j in enumerate(joints)], dtype=np.uint32)
    packed = np.zeros((nw, jmax), dtype=sf._ffi.JOINT_DESC_DTYPE)
    for w, j in enumerate(joints):
        packed[w, :counts[w]] = j[:counts[w]]
    return np.stack(worlds), packed, counts


def test_batch_joints_match_oracle(gpu_api, ref_api):
    """distance / rope joints inside batched worlds (sfp_batch_set_joints), per-world counts"""
    nw, frames = 4, 60
    bodies, joints, counts = rope_worlds(nw)
    batch = sf.Batch(gpu_api, nw, bodies.shape[1])
    batch.upload(bodies)
    batch.set_joints(joints, counts)
    refs = []
    for w in range(nw):
        r = sf.World(ref_api)
        r.upload_bodies(bodies[w])
        r.set_joints(joints[w, :counts[w]])
        refs.append(r)
    for f in range(frames):
        batch.step()
        for r in refs:
            r.step()
    out, st = batch.download(), batch.stats()
    for w in range(nw):
        same(out[w], refs[w].download_bodies())
        rs = refs[w].stats()
        for k in ("n_pairs", "n_contacts", "n_colours", "colour_rounds", "n_joints"):
            assert st[k][w] == rs[k], (w, k, st[k][w], rs[k])
    # ropes hold: no link has fallen to the floor region... (stretch is bounded)
    assert st["n_joints"].sum() == counts.sum()
    # removing the joints again is the joint-free batch
    batch.set_joints(joints[:, :0])
    for r in refs:
        r.set_joints(joints[0, :0])
    for f in range(10):
        batch.step()
        for r in refs:
            r.step()
    out = batch.download()
    for w in range(nw):
        same(out[w], refs[w].download_bodies())


def test_batch_queue_schedule_with_joints(gpu_api):
    """the joint-and-polygon variant of the frame kernel (fewer resident CTAs) through the queue"""
    b4, j4, c4 = rope_worlds(4)
    rep = 90
    bodies, joints, counts = np.tile(b4, (rep, 1)), np.tile(j4, (rep, 1)), np.tile(c4, rep)
    nw = 4 * rep
    outs = []
    for flags in (0, sf.FLAG_BATCH_WHOLE_WORLDS):
        batch = sf.Batch(gpu_api, nw, bodies.shape[1], flags=flags)
        batch.upload(bodies)
        batch.set_joints(joints, counts)
        for _ in range(30):
            batch.step()
        outs.append(batch.download())
        del batch
    same(outs[0].reshape(-1), outs[1].reshape(-1))
    for w in range(4, nw):  # the tiled worlds are copies of the first four
        same(outs[0][w], outs[0][w % 4])


def test_batch_set_joints_rejects_bad_arguments(gpu_api):
    bodies, joints, counts = rope_worlds(2)
    batch = sf.Batch(gpu_api, 2, bodies.shape[1])
    batch.upload(bodies)
    bad = joints.copy()
    bad[1, 0]["body_b"] = bodies.shape[1]  # out of the world
    with pytest.raises(sf.SfpError):
        batch.set_joints(bad, counts)
    with pytest.raises(sf.SfpError):
        batch.set_joints(joints, counts + joints.shape[1])  # more than the slots


def _device_count():
    import ctypes

    try:
        rt = ctypes.CDLL("libcudart.so")
    except OSError:
        import torch

        return torch.cuda.device_count()
    n = ctypes.c_int(0)
    rt.cudaGetDeviceCount(ctypes.byref(n))
    return n.value


@pytest.mark.parametrize("n_blocks", [2, 3])
def test_multi_gpu_entry_equals_single_device_batch(gpu_api, n_blocks):
    """sfp_batch_create_multi: the library spreads the worlds over the devices it is given (one
    SfpBatch + worker thread per block).  Block d runs on GPU d when the box has that many GPUs,
    otherwise the blocks share GPU 0 — the fan-out / join / partition code is the same."""
    nw, bpw, frames = 37, 128, 25   # 37 worlds over 2 / 3 blocks: uneven partition
    bodies = scenes.worlds(nw, bpw)
    ndev = max(1, _device_count())
    devices = [d % ndev for d in range(n_blocks)]
    m = sf.BatchMulti(gpu_api, nw, bpw, devices)
    part = m.partition()
    assert [p[0] for p in part] == devices
    assert part[0][1] == 0 and sum(p[2] for p in part) == nw
    assert all(part[d][1] + part[d][2] == part[d + 1][1] for d in range(n_blocks - 1))
    assert max(p[2] for p in part) - min(p[2] for p in part) <= 1
    one = sf.Batch(gpu_api, nw, bpw)
    m.upload(bodies)
    one.upload(bodies)
    for _ in range(frames):
        m.step()
        one.step()
    m.sync()
    a, b = m.download(), one.download()
    assert a.tobytes() == b.tobytes()
    sa, sb = m.stats(), one.stats()
    for k in ("n_pairs", "n_contacts", "n_colours", "overflow"):
        assert np.array_equal(sa[k], sb[k]), k
    # the per-frame host round trip through the same entry
    out_m = m.step_host(a.reshape(-1))
    out_1 = one.step_host(b.reshape(-1))
    assert out_m.tobytes() == out_1.tobytes()
    assert m.last_step_ms() > 0.0
