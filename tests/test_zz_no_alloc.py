"""sfp_step_small_noalloc, the entry point for hosts that must not allocate while running.

Some users of this library step physics from a thread that may never touch the device
allocator: a game's frame thread (cudaMalloc / cudaFree serialise against the whole device
and would stall the renderer's streams), or an engine that hands out all GPU memory at start-up
and leaves none for a library to claim later.  For them the library carries one fixed ~2.2 MiB
device array and an entry point that does all of its work inside it.

The check: its result must be bit-identical (the tolerance every parity test here uses: zero)
to the ordinary, allocating entry points sfp_batch_create / upload / step / download on the
same world — on two calls in a row, because the array is shared and keeps the previous call's
contents, so the second call is the one that would expose stale state.  The two calls use
different worlds of different sizes and frame counts.

This file sorts last on purpose: a `pytest -x` run has exercised every other test before it.
"""
import numpy as np
import pytest

import starframe_b200 as sf
from starframe_b200 import _ffi, scenes

pytestmark = pytest.mark.gpu
FIELDS = ("x", "y", "c", "s", "vx", "vy", "omega")


def reference_result(api, bodies, frames):
    """the existing (allocating) path: a batch of one world"""
    b = sf.Batch(api, 1, len(bodies))
    b.upload(bodies)
    for _ in range(frames):
        b.step()
    out = b.download()[0]
    st = b.stats()
    b.close()
    return out, st


def assert_bit_identical(a, b):
    for f in FIELDS:
        bad = np.nonzero(a[f].view(np.uint32) != b[f].view(np.uint32))[0]
        assert bad.size == 0, (f, bad[:5], a[f][bad[:5]], b[f][bad[:5]])


def test_no_alloc_entry_matches_batch_entry_twice_in_a_row(gpu_api):
    # call 1: a 512-body pile, long enough to be in contact, to pass the recolouring period
    # (32) and to leave plenty of state behind in the shared array
    world1 = scenes.small_world(3, 512)[0]
    got1 = _ffi.step_small_noalloc(gpu_api, world1, frames=70)
    want1, st1 = reference_result(gpu_api, world1, 70)
    assert st1["n_contacts"][0] > 100  # the comparison is about a world that really collides
    assert_bit_identical(got1, want1)

    # call 2, straight after, on the same array: different world, size and frame count
    world2 = scenes.small_world(11, 200)[0]
    got2 = _ffi.step_small_noalloc(gpu_api, world2, frames=45)
    want2, _ = reference_result(gpu_api, world2, 45)
    assert_bit_identical(got2, want2)

    # and the first world again gives the first answer again: nothing leaked between calls
    again = _ffi.step_small_noalloc(gpu_api, world1, frames=70)
    assert again.tobytes() == got1.tobytes()


def test_no_alloc_entry_refuses_worlds_beyond_its_array(gpu_api):
    too_big = scenes.mixed_pile(1100)[0]
    assert len(too_big) > 1024
    with pytest.raises(sf.SfpError) as e:
        _ffi.step_small_noalloc(gpu_api, too_big, frames=1)
    assert e.value.code == _ffi.SFP_E_CAPACITY
    assert "at most 1024" in str(e.value)
    # a refused call leaves the entry point usable
    ok = scenes.small_world(1, 64)[0]
    got = _ffi.step_small_noalloc(gpu_api, ok, frames=5)
    want, _ = reference_result(gpu_api, ok, 5)
    assert_bit_identical(got, want)


def test_no_alloc_entry_and_batch_entry_overflow_at_the_same_point(gpu_api):
    """The two entry points give a world the same pair capacity (12 n + 256), so a small dense
    world that has more candidate pairs than that is refused by BOTH, and one just under the
    limit is stepped by both with identical bits."""
    def blob(n_side, margin):
        b = scenes.empty_bodies(n_side * n_side)
        k = np.arange(n_side * n_side)
        b["x"] = 0.2 * (k % n_side) + 0.003 * scenes.uniform(7, 1, len(k))
        b["y"] = 0.12 + 0.2 * (k // n_side) + 0.003 * scenes.uniform(7, 2, len(k))
        b["shape"], b["radius"] = 0, 0.09
        return np.concatenate([b, scenes.static_rect(2.0, -0.5, 5.0, 0.5)]), margin

    # ~22 candidate pairs per body with the wide margin: 13 k pairs for 626 bodies > 12 n + 256
    dense, margin = blob(25, 0.26)
    with pytest.raises(sf.SfpError) as e1:
        _ffi.step_small_noalloc(gpu_api, dense, frames=2, aabb_margin=margin)
    assert e1.value.code == _ffi.SFP_E_CAPACITY
    batch = sf.Batch(gpu_api, 1, len(dense), aabb_margin=margin)
    batch.upload(dense)
    batch.step()
    with pytest.raises(sf.SfpError) as e2:
        batch.stats()
    assert e2.value.code == _ffi.SFP_E_CAPACITY
    # the default margin: a handful of pairs per body, both succeed and agree
    loose, _ = blob(25, 0.02)
    got = _ffi.step_small_noalloc(gpu_api, loose, frames=20)
    want, _ = reference_result(gpu_api, loose, 20)
    assert_bit_identical(got, want)
