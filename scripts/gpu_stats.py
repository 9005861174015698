"""Prints per-frame statistics of a workload on the GPU (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from bench import make_scene

name = sys.argv[1] if len(sys.argv) > 1 else "pile100k"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 300
every = int(sys.argv[3]) if len(sys.argv) > 3 else 20
b, j = make_scene(name)
w = sf.World(sf.product_api())
w.upload_bodies(b); w.set_joints(j)
for f in range(frames):
    w.step()
    if f % every == every - 1:
        try:
            st = w.stats()
        except Exception as e:
            print(f, "ERR", e); break
        br = w.last_step_breakdown_ms()
        print(f, {k: st[k] for k in ("n_entries", "n_pairs", "n_contacts", "n_colours", "colour_rounds", "overflow")},
              "pen %.4f ke %.1f" % (st["max_penetration"], st["kinetic_energy"]),
              {k: round(v, 3) for k, v in br.items()}, flush=True)
pairs = w.get_pairs()
deg = np.bincount(np.concatenate([pairs["a"], pairs["b"]]), minlength=len(b))
dyn = (b["flags"] & 1) == 0
print("max dynamic degree", deg[dyn].max(), "mean", deg[dyn].mean(), "top", np.sort(deg[dyn])[-8:])
i = int(np.argmax(np.where(dyn, deg, 0)))
print("body", i, b[i])
