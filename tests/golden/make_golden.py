"""Generates the golden fixtures in this directory.

There are NO reference golden vectors: the reference tree, its tests and its scenes are not
mounted.  These fixtures are produced by the CPU ORACLE (oracle/libsfp_ref.so)
and pin two things: (1) the oracle against accidental change (tests/test_golden.py, CPU tier),
(2) the CUDA path against recorded outputs without the oracle in the loop (GPU tier).

    python tests/golden/make_golden.py        # rewrites tests/golden/*.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import starframe_b200 as sf  # noqa: E402
from oracle import loader  # noqa: E402
from starframe_b200 import scenes  # noqa: E402

def _polys():
    b, j, table = scenes.polygon_pile(500)
    return b, j, table


CASES = {
    "polygons500_f120": (_polys, 120),
    "stack100_f90": (lambda: scenes.stack100(), 90),
    "soup300_f40": (lambda: scenes.random_soup(300, 77, pitch=0.55), 40),
    "pile800_f150": (lambda: scenes.mixed_pile(800), 150),
    "ropes_small_f60": (lambda: scenes.ropes(n_ropes=24, links=20, n_free=150), 60),
    "circles600_f120": (lambda: scenes.circles_container(600, width=4.0), 120),
}


def run(api, bodies, joints, frames, polygons=None):
    w = sf.World(api)
    if polygons is not None:
        w.set_polygons(polygons)
    w.upload_bodies(bodies)
    w.set_joints(joints)
    for _ in range(frames):
        w.step()
    pairs, pc, jc = w.get_colours()
    st = w.stats()
    keys = ("n_entries", "n_pairs", "n_contacts", "n_manifolds", "n_colours", "colour_rounds")
    return dict(state=w.download_bodies(), pairs=pairs, pair_colours=pc, joint_colours=jc,
                contacts=w.get_contacts(), counts=np.array([st[k] for k in keys], dtype=np.int64),
                max_penetration=np.float32(st["max_penetration"]))


def main():
    api = loader.api()
    for name, (make, frames) in CASES.items():
        made = make()
        bodies, joints = made[0], made[1]
        polygons = made[2] if len(made) > 2 else np.zeros(0, dtype=sf._ffi.POLYGON_DTYPE)
        out = run(api, bodies, joints, frames, polygons)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), bodies=bodies, joints=joints,
                            polygons=polygons, frames=np.int64(frames), **out)
        print(name, len(bodies), "bodies", len(out["pairs"]), "pairs", out["counts"].tolist())


if __name__ == "__main__":
    main()
