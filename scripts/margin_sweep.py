import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from starframe_b200 import scenes
b, j = scenes.mixed_pile(100_000)
for margin in (0.02, 0.015, 0.011):
    w = sf.World(sf.product_api(), aabb_margin=margin)
    w.upload_bodies(b)
    for f in range(300): w.step()
    t = 0
    for f in range(20):
        w.step(); w.sync(); t += w.last_step_ms()
    st = w.stats()
    pairs, pc, _ = w.get_colours()
    print(margin, "ms", round(t / 20, 3), {k: st[k] for k in ("n_pairs", "n_contacts", "n_manifolds", "n_colours", "max_penetration")}, "big colours", int((np.bincount(pc) > 512).sum()))
