"""Long run of GPU against oracle on small scenes, compared every 50 frames (development aid):
  python scripts/soak.py [frames]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from starframe_b200 import scenes
from oracle import loader
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
cases = {"pile3000": scenes.mixed_pile(3000), "ropes": scenes.ropes(n_ropes=40, links=32, n_free=300),
         "circles2000": scenes.circles_container(2000), "polys1500": scenes.polygon_pile(1500)}
for name, scene in cases.items():
    g, r = sf.World(sf.product_api()), sf.World(loader.api())
    for w in (g, r):
        if len(scene) > 2: w.set_polygons(scene[2])
        w.upload_bodies(scene[0]); w.set_joints(scene[1])
    bad = 0
    for f in range(frames):
        g.step(); r.step()
        if (f + 1) % 50 == 0:
            a, c = g.download_bodies(), r.download_bodies()
            bad += sum(int(np.count_nonzero(a[k].view(np.uint32) != c[k].view(np.uint32))) for k in ("x", "y", "c", "s", "vx", "vy", "omega"))
            if not np.array_equal(g.get_pairs(), r.get_pairs()): bad += 1
    print(name, frames, "frames: words that differ", bad, "colours", g.stats()["n_colours"], r.stats()["n_colours"], flush=True)
# batched worlds, queue schedule, against the oracle per world
nw, bpw = 400, 64
bodies = scenes.worlds(nw, bpw)
b = sf.Batch(sf.product_api(), nw, bpw); b.upload(bodies)
refs = []
for w in range(0, nw, 57):
    r = sf.World(loader.api()); r.upload_bodies(bodies[w]); refs.append((w, r))
bad = 0
for f in range(frames):
    b.step()
    for _, r in refs: r.step()
    if (f + 1) % 100 == 0:
        out = b.download()
        for w, r in refs:
            c = r.download_bodies()
            bad += sum(int(np.count_nonzero(out[w][k].view(np.uint32) != c[k].view(np.uint32))) for k in ("x", "y", "c", "s", "vx", "vy", "omega"))
print("batch", nw, "worlds", frames, "frames: words that differ", bad)
